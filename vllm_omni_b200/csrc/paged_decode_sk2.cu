// Paged decode attention for sm_100a — stream-K, second generation (the default): ONE kernel per
// layer, no reduce launch, a plan computed once per step.
//
// Same tile stream as paged_decode_streamk.cu: all 16-token KV tiles of the step, ordered
// (sequence, kv head [, head sub-group], tile), cut into NW = SMs x W equal contiguous chunks, one
// per WARP of a persistent grid; each warp streams its chunk through a private ring of TMA stages
// and runs the tensor-core math on the swizzled tiles (sk_math.cuh). What changed, and why
// (profiles/README.md, round 1: 9.0 us fixed cost + 4.5 us reduce launch against 5.7 us for the
// TRT-LLM-gen kernel):
//
//  * The plan is not rebuilt by every CTA. sk2_plan_kernel runs ONCE PER STEP (the block table and
//    sequence lengths are the same for all layers): tiles prefix, per-warp start position and the
//    warp's first 32 block ids. A warp's prologue is one round of independent loads (descriptor +
//    ids) and then its first TMA — no __syncthreads, no binary search, no dependent id load.
//  * Early issue under programmatic dependent launch: the block table, sequence lengths and plan
//    are not written by any kernel of the step, and the only KV bytes a still-running predecessor
//    (the cache write of the same layer) can touch are the new token's slot, i.e. the LAST tile of
//    a sequence. With B200ATTN_DECODE_EARLY_KV_READ a warp therefore fills its ring before
//    griddepcontrol.wait and only waits before (a) loading q, (b) issuing the last tile of a
//    sequence, (c) its first global store. In a chain write -> decode -> write -> decode the next
//    layer's first stages stream while the previous layer's stragglers finish.
//  * SMs do not stream at equal rates under a saturated HBM (a property of where an SM sits: the same SM
//    pairs finish 30 % later than the fastest in every launch), so the chunk group a CTA runs is the one of
//    the SM it sits on, sized by the plan kernel in proportion to the rate that SM achieved before.
//  * The separate log-sum-exp reduce kernel (sk2_reduce_kernel) is plan-driven too: which chunks cut an item
//    is in the plan, so its CTAs (launched early, programmatic dependent launch) have their piece lists
//    before the decode kernel ends and need one round of loads after it. (Merging inside the decode kernel
//    — CTA-cooperative, last-arriver across CTAs — was built and measured: its chain of dependent L2 round
//    trips at the very end of the kernel cost 8 us against 3.4 us for the pre-launched reduce kernel.)
//
// Replaces the decode branch of the reference's stock backends
// (vllm/v1/attention/backends/flashinfer.py:1694-1822, flash_attn.py:758-819) and their split-KV
// combine (vllm_flash_attn flash_fwd_combine; vllm/v1/attention/ops/triton_merge_attn_states.py).
#include "kernels.h"
#include "sk_math.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------
// Plan layout (int32 words): header | cstart[nw + 1] | desc[nw][4] | ids[nw][32] | cum[batch + 1] | item_w[items][2]
//   cstart[c]  first stream position (tile x head-group) of chunk c; chunk c = CTA group c / W, warp c % W
//   desc[c]    {sequence, head-group, tile, tiles of the sequence} at cstart[c]
//   ids[c]     the 32 block ids of the page window chunk c starts in
//   cum[b]     exclusive prefix of tiles per sequence
//   item_w[i]  first and last chunk that hold tiles of item i = sequence * hs_count + head-group (reduce kernel)
// ------------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ int64_t sk2_plan_cstart_off() { return kSk2PlanHdrInts; }
__host__ __device__ __forceinline__ int64_t sk2_plan_desc_off(int nw) { return kSk2PlanHdrInts + ((nw + 1 + 3) & ~3); }
__host__ __device__ __forceinline__ int64_t sk2_plan_ids_off(int nw) { return sk2_plan_desc_off(nw) + 4ll * nw; }
__host__ __device__ __forceinline__ int64_t sk2_plan_cum_off(int nw) { return sk2_plan_desc_off(nw) + 36ll * nw; }

__host__ __device__ __forceinline__ int64_t sk2_plan_itemw_off(int nw, int batch) {
  return sk2_plan_cum_off(nw) + ((batch + 1 + 3) & ~3);
}

size_t sk2_plan_ints(int batch, int nw, int hs_count) {
  return static_cast<size_t>(sk2_plan_itemw_off(nw, batch)) + 2ull * batch * hs_count;
}
size_t sk2_ctrl_bytes() { return sizeof(Sk2Ctrl); }

struct Sk2Pos {
  int b;     // sequence
  int hs;    // kv head * subgroups + head sub-group
  int tile;  // 16-token tile inside the sequence
  int n;     // tiles of sequence b
};

// Chunk that holds stream position t: the last c in [0, nw) with cstart[c] <= t (empty chunks share their
// successor's start, so the last one is the chunk that really holds t).
__device__ __forceinline__ int sk2_chunk_of(const int32_t* cstart, int nw, int t) {
  int lo = 0, hi = nw;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (cstart[mid] <= t) lo = mid; else hi = mid;
  }
  return lo;
}

// One CTA, once per step.
//  Phase 0: SM weights. SMs do not stream at equal rates under a saturated HBM — a property of where an SM
//           sits (the same SM pairs finish 30 % later than the fastest in every launch, profiles/README.md) —
//           so CTA group g, which is run by the CTA on SM g, gets a share of the stream proportional to the
//           rate that SM achieved in the previous launches (tiles / cycles, recorded by the decode kernel).
//           The weights are frozen after `calib_max` updates: from then on a launch is a deterministic
//           function of its inputs.
//  Phase 1: exclusive prefix of tiles per sequence. Phase 2: chunk starts (group range cut into W equal
//           chunks) and their positions. Phase 3: the 32 block ids of the page window each chunk starts in.
constexpr int kSk2PlanSmemBatch = 4096;  // cum[] of batches up to this size is searched in shared memory

__device__ __forceinline__ float sk2_block_sum(float v, float* s_red, int tid) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  __syncthreads();  // s_red may still be read from the previous call
  if ((tid & 31) == 0) s_red[tid >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int w = 0; w < 32; ++w) t += s_red[w];
  return t;
}

__global__ void __launch_bounds__(1024)
sk2_plan_kernel(const int32_t* __restrict__ seq_lens, int batch, int hs_count, int groups, int warps,
                const int32_t* __restrict__ block_table, int64_t bt_stride, int tiles_per_page, int tpp_shift,
                int32_t* __restrict__ plan, Sk2Ctrl* __restrict__ ctrl, int calib_max) {
  __shared__ int s_warp[32];
  __shared__ float s_red[32];
  __shared__ int s_carry;
  __shared__ float s_w[kSk2MaxGroups];
  __shared__ int s_gstart[kSk2MaxGroups + 1];
  __shared__ int s_cum[kSk2PlanSmemBatch + 1];
  __shared__ int s_cstart[kSk2MaxGroups * 12 + 1];
  const int nw = groups * warps;
  int32_t* cstart = plan + sk2_plan_cstart_off();
  int32_t* cum = plan + sk2_plan_cum_off(nw);
  int4* desc = reinterpret_cast<int4*>(plan + sk2_plan_desc_off(nw));
  int32_t* ids = plan + sk2_plan_ids_off(nw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  // ---- phase 0: weights (thread g <-> SM g)
  {
    const bool have = tid < groups;
    const int tiles = have ? ctrl->stat_tiles[tid] : 0, cyc = have ? ctrl->stat_cycles[tid] : 0;
    const float w_raw = have ? ctrl->weight[tid] : 0.f;
    const float rate = (tiles >= kSk2CalibMinTiles && cyc > 0) ? static_cast<float>(tiles) / static_cast<float>(cyc) : 0.f;
    const int missing = __syncthreads_or(have && rate == 0.f);  // some SM has no (or too little) history
    const bool upd = !missing && ctrl->calib_updates < calib_max;
    const float rsum = sk2_block_sum(rate, s_red, tid), wsum = sk2_block_sum(w_raw, s_red, tid);
    const float w_old = wsum > 0.f ? w_raw / wsum : 1.f / groups;
    float w = upd ? (1.f - kSk2CalibGain) * w_old + kSk2CalibGain * rate / rsum : w_old;
    if (!have) w = 0.f;
    const float norm = sk2_block_sum(w, s_red, tid);
    w /= norm;
    if (have) {
      s_w[tid] = w;
      if (upd) {
        ctrl->weight[tid] = w;
        ctrl->stat_tiles[tid] = 0;  // a history is used once
        ctrl->stat_cycles[tid] = 0;
      }
    }
    __syncthreads();
    if (tid == 0 && upd) ctrl->calib_updates += 1;
  }
  // ---- phase 1
  for (int base = 0; base < batch; base += 1024) {
    const int i = base + tid;
    const int s = i < batch ? max(0, seq_lens[i]) : 0;
    const int t = (s + kTileTokens - 1) / kTileTokens;
    int incl = t;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int wbase = 0;
    for (int w = 0; w < warp; ++w) wbase += s_warp[w];
    const int carry = s_carry;
    if (i < batch) {
      cum[i] = carry + wbase + incl - t;
      if (i < kSk2PlanSmemBatch) s_cum[i] = carry + wbase + incl - t;
    }
    __syncthreads();
    if (tid == 1023) s_carry = carry + wbase + incl;
    __syncthreads();
  }
  const int tiles = s_carry;
  if (static_cast<int64_t>(tiles) * hs_count > 0x7fffffffll) __trap();  // stream positions are 32-bit
  const int total = tiles * hs_count;
  const bool cum_in_smem = batch <= kSk2PlanSmemBatch;
  const int32_t* cum_r = cum_in_smem ? s_cum : cum;
  if (tid == 0) {
    cum[batch] = tiles;
    if (cum_in_smem) s_cum[batch] = tiles;
    plan[0] = total;
    plan[1] = groups;
    plan[2] = warps;
    plan[3] = nw;
    plan[4] = batch;
    plan[5] = hs_count;
    plan[6] = tpp_shift;
    plan[7] = kSk2PlanMagic;
    // group ranges: cumulative weights; every group gets at least one position while there are any, so the
    // non-empty groups are a prefix of the grid
    double acc = 0.0;
    s_gstart[0] = 0;
    for (int g = 0; g < groups; ++g) {
      acc += s_w[g];
      int e = static_cast<int>(acc * total + 0.5);
      e = min(max(e, s_gstart[g] + 1), total);
      s_gstart[g + 1] = g + 1 == groups ? total : e;
    }
  }
  __syncthreads();  // cum[] and s_gstart[] are read below
  // ---- phase 2
  for (int c = tid; c <= nw; c += 1024) {
    const int g = c / warps, w = c % warps;
    // a group's range is cut into `warps` chunks, the first (len % warps) one position longer: the non-empty
    // chunks are a prefix of the group as well
    const int gs = s_gstart[min(g, groups)], ge = s_gstart[min(g + 1, groups)];
    const int q = (ge - gs) / warps, r = (ge - gs) % warps;
    const int g0 = c == nw ? total : gs + w * q + min(w, r);
    cstart[c] = g0;
    s_cstart[c] = g0;
    if (c == nw) break;
    int4 d = make_int4(batch, 0, 0, 1);
    if (g0 < total) {
      const int key = g0 / hs_count;  // cum[b] <= key < cum[b+1]
      int lo = 0, hi = batch;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (cum_r[mid] <= key) lo = mid; else hi = mid;
      }
      const int n = cum_r[lo + 1] - cum_r[lo];
      const int rem = g0 - cum_r[lo] * hs_count;
      d = make_int4(lo, rem / n, rem % n, n);
    }
    desc[c] = d;
  }
  __syncthreads();
  // ---- phase 3
  for (int idx = tid; idx < nw * 32; idx += 1024) {
    const int4 d = desc[idx >> 5];
    const int page0 = (d.z >> tpp_shift) & ~31;
    const int pages = (d.w + tiles_per_page - 1) >> tpp_shift;
    const int pg = page0 + (idx & 31);
    ids[idx] = (d.x < batch && pg < pages) ? block_table[static_cast<int64_t>(d.x) * bt_stride + pg] : 0;
  }
  // ---- phase 4: the chunks that cut each item
  int2* item_w = reinterpret_cast<int2*>(plan + sk2_plan_itemw_off(nw, batch));
  for (int it = tid; it < batch * hs_count; it += 1024) {
    const int b = it / hs_count, hs = it % hs_count;
    const int c0 = cum_r[b], n = cum_r[b + 1] - c0;
    int2 w = make_int2(0, -1);
    if (n > 0) {
      const int t0 = c0 * hs_count + hs * n;
      w = make_int2(sk2_chunk_of(s_cstart, nw, t0), sk2_chunk_of(s_cstart, nw, t0 + n - 1));
    }
    item_w[it] = w;
  }
}

cudaError_t launch_sk2_plan(const int32_t* seq_lens, int batch, int hs_count, int groups, const int32_t* block_table,
                            int64_t bt_stride, int tiles_per_page, int tpp_shift, int32_t* plan, Sk2Ctrl* ctrl,
                            int calib_max, cudaStream_t stream) {
  sk2_plan_kernel<<<1, 1024, 0, stream>>>(seq_lens, batch, hs_count, groups, decode_sk2_warps(), block_table, bt_stride,
                                          tiles_per_page, tpp_shift, plan, ctrl, calib_max);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// The separate log-sum-exp reduce kernel: merge of the partials left by chunk boundaries (the
// merge_attn_states formula). One CTA of 8 warps per (sequence, kv head, head sub-group) item; a partial
// is one contiguous 4 KiB block (8 heads x 128 floats) plus 8 LSEs, so warp w folds the pieces
// k = w, w + 8, ... of all 8 heads with 8 independent 512 B loads in flight per piece, and the 8 warps are
// folded through shared memory. The item's piece list comes from the plan (first / last chunk, chunk
// starts), read BEFORE griddepcontrol.wait. Items finished by a single warp were already written by it;
// rows of empty sequences are zeroed here.
// ------------------------------------------------------------------------------------------------
constexpr int kSk2ReduceWarps = 8;

__global__ void __launch_bounds__(kSk2ReduceWarps * 32)
sk2_reduce_kernel(__nv_bfloat16* __restrict__ out, int64_t out_stride_tok, float* __restrict__ lse, int batch,
                  int num_q_heads, int group, int subgroups, int num_kv_heads, int nw,
                  const float* __restrict__ part_o, const float* __restrict__ part_lse,
                  const int32_t* __restrict__ plan) {
  __shared__ float s_acc[kSk2ReduceWarps][kHeadsPerItem][kHeadDim];
  __shared__ float s_max[kSk2ReduceWarps][kHeadsPerItem];
  __shared__ float s_den[kSk2ReduceWarps][kHeadsPerItem];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int hs_count = num_kv_heads * subgroups;
  const int item = blockIdx.x;
  const int b = item / hs_count, hs = item % hs_count;
  const int kvh = hs / subgroups, sub = hs % subgroups;
  const int heads_here = min(kHeadsPerItem, group - sub * kHeadsPerItem);
  const int h_first = kvh * group + sub * kHeadsPerItem;  // first query head of the item
  // the plan is not written by the kernels this launch depends on: read it while they still run
  const int32_t* cstart = plan + sk2_plan_cstart_off();
  const int2 w = __ldg(reinterpret_cast<const int2*>(plan + sk2_plan_itemw_off(nw, batch)) + item);
  const int pieces = w.y - w.x + 1;  // chunks from the first to the last that hold tiles of the item
  // which of this warp's pieces really exist (a group's trailing chunks may be empty): lane i <-> piece warp + 8 i
  bool mine = false;
  {
    const int k = warp + kSk2ReduceWarps * lane;
    if (k < pieces) mine = __ldg(cstart + w.x + k + 1) > __ldg(cstart + w.x + k);
  }
  const uint32_t live = __ballot_sync(0xffffffffu, mine);
  griddep_wait();
  griddep_launch_dependents();
  if (pieces <= 0) {  // empty sequence: zero rows, LSE = -inf
    if (warp < heads_here) {
      const int h = h_first + warp;
      *reinterpret_cast<uint2*>(out + static_cast<int64_t>(b) * out_stride_tok + static_cast<int64_t>(h) * kHeadDim +
                                lane * 4) = make_uint2(0u, 0u);
      if (lse != nullptr && lane == 0) lse[static_cast<int64_t>(b) * num_q_heads + h] = -CUDART_INF_F;
    }
    return;
  }
  if (pieces == 1) return;  // finished by one warp

  float4 acc[kHeadsPerItem];
  float den[kHeadsPerItem], m[kHeadsPerItem];
#pragma unroll
  for (int hh = 0; hh < kHeadsPerItem; ++hh) {
    acc[hh] = make_float4(0.f, 0.f, 0.f, 0.f);
    den[hh] = 0.f;
    m[hh] = -CUDART_INF_F;
  }
  for (int i = 0; warp + kSk2ReduceWarps * i < pieces; ++i) {
    const int k = warp + kSk2ReduceWarps * i;
    const bool on = i < 32 ? ((live >> i) & 1u) != 0
                           : __ldg(cstart + w.x + k + 1) > __ldg(cstart + w.x + k);  // (more than 256 pieces: rare)
    if (!on) continue;
    const int64_t r0 = (2ll * (w.x + k) + (k == 0 ? 1 : 0)) * kHeadsPerItem;
    const float ls = lane < kHeadsPerItem ? part_lse[r0 + lane] : 0.f;
    float4 o[kHeadsPerItem];
#pragma unroll
    for (int hh = 0; hh < kHeadsPerItem; ++hh)
      o[hh] = *reinterpret_cast<const float4*>(part_o + (r0 + hh) * kHeadDim + lane * 4);
#pragma unroll
    for (int hh = 0; hh < kHeadsPerItem; ++hh) {
      const float l = __shfl_sync(0xffffffffu, ls, hh);
      if (l == -CUDART_INF_F) continue;  // empty partial
      const float mn = fmaxf(m[hh], l);
      const float a = fast_exp2((m[hh] - mn) * kLog2e);  // 0 when m = -inf
      const float wgt = fast_exp2((l - mn) * kLog2e);
      acc[hh].x = fmaf(o[hh].x, wgt, acc[hh].x * a);
      acc[hh].y = fmaf(o[hh].y, wgt, acc[hh].y * a);
      acc[hh].z = fmaf(o[hh].z, wgt, acc[hh].z * a);
      acc[hh].w = fmaf(o[hh].w, wgt, acc[hh].w * a);
      den[hh] = fmaf(den[hh], a, wgt);
      m[hh] = mn;
    }
  }
#pragma unroll
  for (int hh = 0; hh < kHeadsPerItem; ++hh) {
    *reinterpret_cast<float4*>(&s_acc[warp][hh][lane * 4]) = acc[hh];
    if (lane == 0) {
      s_den[warp][hh] = den[hh];
      s_max[warp][hh] = m[hh];
    }
  }
  __syncthreads();
  // ---- fold the 8 warps: warp = head, lane = 4 output columns
  if (warp < heads_here) {
    float mfin = s_max[0][warp];
#pragma unroll
    for (int ww = 1; ww < kSk2ReduceWarps; ++ww) mfin = fmaxf(mfin, s_max[ww][warp]);
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    float dsum = 0.f;
#pragma unroll
    for (int ww = 0; ww < kSk2ReduceWarps; ++ww) {
      const float mw = s_max[ww][warp];
      const float sc_w = (mw == -CUDART_INF_F) ? 0.f : fast_exp2((mw - mfin) * kLog2e);
      const float4 v = *reinterpret_cast<const float4*>(&s_acc[ww][warp][lane * 4]);
      a.x = fmaf(v.x, sc_w, a.x);
      a.y = fmaf(v.y, sc_w, a.y);
      a.z = fmaf(v.z, sc_w, a.z);
      a.w = fmaf(v.w, sc_w, a.w);
      dsum = fmaf(s_den[ww][warp], sc_w, dsum);
    }
    const float inv = dsum > 0.f ? 1.f / dsum : 0.f;
    const int h = h_first + warp;
    uint2 pk;
    pk.x = pack_bf16x2(a.x * inv, a.y * inv);
    pk.y = pack_bf16x2(a.z * inv, a.w * inv);
    *reinterpret_cast<uint2*>(out + static_cast<int64_t>(b) * out_stride_tok + static_cast<int64_t>(h) * kHeadDim +
                              lane * 4) = pk;
    if (lse != nullptr && lane == 0)
      lse[static_cast<int64_t>(b) * num_q_heads + h] = dsum > 0.f ? mfin + logf(dsum) : -CUDART_INF_F;
  }
}

template <int W, int D, int SB>
struct Sk2Smem {
  alignas(1024) uint8_t ring[W][D][SB];  // SB = bytes of one stage (K tile + V tile)
  alignas(8) uint64_t bar[W][D];         // "stage landed" mbarriers, one per ring slot
  long long t0;                          // clock64 at CTA start (the SM's rate is recorded at the end)
  int group;                             // chunk group this CTA claimed
};

__device__ __forceinline__ void sk2_advance(const int32_t* cum, int batch, int hs_count, Sk2Pos& it) {
  if (++it.tile < it.n) return;
  it.tile = 0;
  if (++it.hs < hs_count) return;
  it.hs = 0;
  int c0 = __ldg(cum + it.b + 1), c1 = c0;
  do {
    ++it.b;
    c0 = c1;
    c1 = it.b < batch ? __ldg(cum + it.b + 1) : c0 + 1;
  } while (it.b < batch && c1 == c0);
  it.n = c1 - c0;
}

__device__ __forceinline__ int sk2_load_ids(const DecodeParams& p, int b, int n_tiles, int page0, int lane) {
  const int pages = (n_tiles + p.tiles_per_page - 1) >> p.tpp_shift;
  const int pg = page0 + lane;
  return (b < p.batch && pg < pages) ? __ldg(p.block_table + static_cast<int64_t>(b) * p.bt_stride + pg) : 0;
}

__device__ __forceinline__ uint64_t sk2_globaltimer() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

template <int W, int D, bool FP8>
__global__ void __launch_bounds__(W * 32, 1)
paged_decode_sk2_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                        const __grid_constant__ DecodeParams p) {
  constexpr int SB = FP8 ? kKvTileBytesFp8 : kKvTileBytesBf16;  // stage bytes
  constexpr int VOFF = SB / 2;                                   // V tile behind the K tile
  extern __shared__ uint8_t smem_raw[];
  Sk2Smem<W, D, SB>& sm =
      *reinterpret_cast<Sk2Smem<W, D, SB>*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nw = gridDim.x * W;
  // ---- which chunk group this CTA runs: the one of the SM it sits on (the plan sized group g for SM g), claimed
  // with one atomic so that every group is run exactly once whatever the block scheduler did (two CTAs on one
  // SM one after the other, an SM id beyond the grid): a CTA whose own group is taken scans for a free one.
  // Claim words of two consecutive launches (which may overlap under programmatic dependent launch) are kept
  // apart by the launch parity; the last CTA to finish re-arms the words of its launch.
  Sk2Ctrl* const ctrl = p.sk2_ctrl;
  const int parity = (p.flags & kSk2FlagParity) ? 1 : 0;
  const int G = static_cast<int>(gridDim.x);
  uint32_t smid;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  const int g_spec = (p.flags & kSk2FlagDebugInterleave) ? static_cast<int>(blockIdx.x)
                                                         : static_cast<int>(smid % static_cast<uint32_t>(G));
  if (threadIdx.x == 0) {
    int g = g_spec;
    if (!(p.flags & kSk2FlagDebugInterleave) && atomicCAS(&ctrl->claimed[parity][g], 0, 1) != 0) {
      for (int k = 1; k <= G; ++k) {
        const int gg = (g + k) % G;
        if (atomicCAS(&ctrl->claimed[parity][gg], 0, 1) == 0) {
          g = gg;
          break;
        }
      }
    }
    sm.group = g;
  }
  bool waited = false;
  // griddepcontrol.wait; the SM's rate (recorded for the plan kernel) is measured from here, so the time a CTA
  // that started early spends waiting for the previous kernel does not count against its SM
  auto wait_for_prior = [&]() {
    griddep_wait();
    waited = true;
    if (threadIdx.x == 0) sm.t0 = clock64();
  };
  if (p.flags & kSk2FlagWaitFirst) wait_for_prior();  // the plan was written by the kernel right before this one
  // ---- the warp's work: one round of independent loads, issued for the group the CTA will almost always get
  // (its own SM's) while thread 0's claim is in flight
  const int32_t* plan = p.sk2_plan;
  const int4 hdr = __ldg(reinterpret_cast<const int4*>(plan));
  const int4 hdr2 = __ldg(reinterpret_cast<const int4*>(plan) + 1);
  // chunk of this warp; debug flag: chunks interleaved over the CTAs (timeline experiments only: the merge is skipped)
  int gw = (p.flags & kSk2FlagDebugInterleave) ? warp * G + static_cast<int>(blockIdx.x) : g_spec * W + warp;
  int4 d4 = __ldg(reinterpret_cast<const int4*>(plan + sk2_plan_desc_off(nw)) + gw);
  int g_begin = __ldg(plan + sk2_plan_cstart_off() + gw), g_end = __ldg(plan + sk2_plan_cstart_off() + gw + 1);
  int ids = __ldg(plan + sk2_plan_ids_off(nw) + gw * 32 + lane);
  const int32_t* cum = plan + sk2_plan_cum_off(nw);
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < D; ++s) mbar_init(&sm.bar[warp][s], 1);
    fence_barrier_init();
    if (warp == 0) {
      tma_prefetch_desc(&tm_k);
      tma_prefetch_desc(&tm_v);
    }
  }
  __syncthreads();
  const int group = sm.group;
  if (group != g_spec) {  // rare: this CTA runs another SM's group
    gw = group * W + warp;
    d4 = __ldg(reinterpret_cast<const int4*>(plan + sk2_plan_desc_off(nw)) + gw);
    g_begin = __ldg(plan + sk2_plan_cstart_off() + gw);
    g_end = __ldg(plan + sk2_plan_cstart_off() + gw + 1);
    ids = __ldg(plan + sk2_plan_ids_off(nw) + gw * 32 + lane);
  }
  if (p.trace != nullptr && lane == 0) {  // debug: entry stamp, SM id
    p.trace[gw * 8 + 0] = sk2_globaltimer();
    p.trace[gw * 8 + 4] = smid;
  }
  // the next kernel of the chain may start its own prologue (it cannot get an SM before this CTA exits)
  griddep_launch_dependents();

  const int hs_count = p.num_kv_heads * p.subgroups;
  if (hdr.y != static_cast<int>(gridDim.x) || hdr.z != W || hdr.w != nw || hdr2.x != p.batch || hdr2.y != hs_count ||
      hdr2.z != p.tpp_shift || hdr2.w != kSk2PlanMagic)
    __trap();  // the plan in the workspace was made for another launch shape: fail loudly
  const int cnt = g_end - g_begin;
  const bool early = (p.flags & kSk2FlagEarlyKvRead) != 0;

  const SkLane sl = sk_make_lane(lane);
  const uint32_t ring0 = smem_u32(sm.ring[warp][0]);
  const uint32_t bar0 = smem_u32(&sm.bar[warp][0]);
  const uint64_t pol = l2_policy_evict_first();  // KV pages are streamed once per step

  // prefetch-front state
  Sk2Pos pf{d4.x, d4.y, d4.z, d4.w};
  int pf_left = cnt;
  int pf_page0 = (pf.tile >> p.tpp_shift) & ~31, pf_kvh = 0;
  int pf_items_ahead = 0;   // item starts prepared by the front and not yet taken by the math (0 or 1)
  bool pf_fresh = true;     // the front stands on a tile whose item it has not prepared yet
  bool first_item = true;   // ids of the first item came with the plan
  bool q_deferred = false;  // the prepared item's q has not been loaded (griddepcontrol.wait pending)
  int it_b = 0, it_hs = 0;  // the item the front prepared last (the front itself may have moved on)
  int pslot = 0, pending = 0;
  // math state
  Sk2Pos cs = pf;
  int cs_left = cnt, cslot = 0;
  uint32_t cphase = 0;  // parity of the current pass over the ring
  int seg_start = cs.tile;
  bool have_item = false;
  uint32_t qf[8][2], qf_n[8][2];
#pragma unroll
  for (int kk = 0; kk < 8; ++kk) qf[kk][0] = qf[kk][1] = qf_n[kk][0] = qf_n[kk][1] = 0u;
  float scale = p.scale_log2, scale_n = p.scale_log2;
  int slen = 0, slen_n = 0;
  WarpState st;
  state_reset(st);

  auto load_q_next = [&]() {  // q (and the K scale) of the item the front prepared last
    WorkItem w;
    w.b = it_b;
    w.kvh = it_hs / p.subgroups;
    w.sub = it_hs % p.subgroups;
    if (FP8) {
      sk_load_q_f16(p, w.b, w.kvh, w.sub, lane, qf_n);
      scale_n = p.scale_log2 * __ldg(p.k_scale + w.kvh * p.scale_stride);
    } else {
      load_q_frags(p, w, lane, qf_n);
    }
  };

  while (cs_left > 0) {
    // ================= keep the ring full =================
    while (pending < D && pf_left > 0) {
      if (pf_fresh) {
        // first tile (of this chunk) of an item: qf_n holds one prepared item, so the front never runs two
        // item starts ahead of the math
        if (pf_items_ahead == 1) break;
        it_b = pf.b;
        it_hs = pf.hs;
        pf_kvh = pf.hs / p.subgroups;
        slen_n = max(0, __ldg(p.seq_lens + pf.b));
        if (!first_item) {
          pf_page0 = 0;
          ids = sk2_load_ids(p, pf.b, pf.n, 0, lane);
        }
        first_item = false;
        if (waited) load_q_next(); else q_deferred = true;
        pf_items_ahead = 1;
        pf_fresh = false;
      }
      if (!waited && (!early || pf.tile + 1 == pf.n)) {
        // without the early-read contract every KV byte may still be in flight; with it only the tile that
        // holds the new token's slot
        wait_for_prior();
        if (q_deferred) {
          load_q_next();
          q_deferred = false;
        }
      }
      const int page = pf.tile >> p.tpp_shift;
      if (page >= pf_page0 + 32) {
        pf_page0 = page & ~31;
        ids = sk2_load_ids(p, pf.b, pf.n, pf_page0, lane);
      }
      const int blk = __shfl_sync(0xffffffffu, ids, page - pf_page0);
      if (lane == 0) {
        const uint32_t dst = ring0 + pslot * SB, bar = bar0 + pslot * 8;
        const int slot0 = (pf.tile & p.tpp_mask) * kTileTokens;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(SB) : "memory");
        if (FP8) {
          tma_load_4d_s(dst, &tm_k, bar, 0, slot0, pf_kvh, blk, pol);
          tma_load_4d_s(dst + VOFF, &tm_v, bar, 0, slot0, pf_kvh, blk, pol);
        } else {
          tma_load_5d(dst, &tm_k, bar, 0, slot0, 0, pf_kvh, blk, pol);
          tma_load_5d(dst + VOFF, &tm_v, bar, 0, slot0, 0, pf_kvh, blk, pol);
        }
      }
      ++pending;
      --pf_left;
      pslot = (pslot + 1 == D) ? 0 : pslot + 1;
      sk2_advance(cum, p.batch, hs_count, pf);
      pf_fresh = (pf.tile == 0);
    }
    if (q_deferred) {  // the first stages are in flight: now wait for the previous kernel and fetch q
      wait_for_prior();
      load_q_next();
      q_deferred = false;
    }
    if (!have_item) {  // take the item the front prepared
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        qf[kk][0] = qf_n[kk][0];
        qf[kk][1] = qf_n[kk][1];
      }
      scale = scale_n;
      slen = slen_n;
      pf_items_ahead = 0;
      have_item = true;
      continue;  // the front may now prepare the item after this one
    }

    // ================= math on the oldest stage =================
    mbar_wait(&sm.bar[warp][cslot], cphase);
    if (p.trace != nullptr && cs_left == cnt && lane == 0) p.trace[gw * 8 + 1] = sk2_globaltimer();  // first stage landed
    const bool item_done = (cs.tile + 1 == cs.n);
    {
      const int valid = item_done ? slen - cs.tile * kTileTokens : kTileTokens;
      if (FP8)
        sk_consume_fp8<true>(ring0 + cslot * SB, sm.ring[warp][cslot], valid, qf, sl, scale, st, lane);
      else
        sk_consume<true>(ring0 + cslot * SB, sm.ring[warp][cslot], valid, qf, sl, scale, st, lane);
      if (valid < kTileTokens) fence_proxy_async_smem();  // zeroed V rows vs. the next TMA write
    }
    __syncwarp();  // every lane has read the slot before lane 0 refills it
    --pending;
    --cs_left;
    if (item_done || cs_left == 0) {
      // the slot just drained is free: use it to transpose the fragments
      const bool whole = item_done && seg_start == 0;
      const int piece = gw * 2 + (seg_start > 0 ? 0 : 1);
      sk_store<FP8>(p, st, cs.b, cs.hs, whole, piece, reinterpret_cast<float*>(sm.ring[warp][cslot]), lane);
      fence_proxy_async_smem();  // the slot was written through the generic proxy; TMA refills it next
      __syncwarp();
      state_reset(st);
      seg_start = 0;
      have_item = false;
    }
    if (++cslot == D) {
      cslot = 0;
      cphase ^= 1u;
    }
    sk2_advance(cum, p.batch, hs_count, cs);
  }

  // every warp waits before it exits, so a completed grid implies its predecessors completed
  if (!waited) wait_for_prior();
  if (p.trace != nullptr && lane == 0) {  // debug: end of this warp's streaming
    const uint64_t t_end = sk2_globaltimer();
    if (cnt == 0) p.trace[gw * 8 + 1] = t_end;
    p.trace[gw * 8 + 2] = t_end;
    p.trace[gw * 8 + 3] = t_end;
  }
  if (p.flags & kSk2FlagDebugInterleave) return;
  __syncthreads();
  if (threadIdx.x == 0) {
    // what this SM achieved (the plan kernel turns it into the SM's share of the next steps' streams) ...
    if (smid < static_cast<uint32_t>(kSk2MaxGroups)) {
      const int32_t* cstart = plan + sk2_plan_cstart_off();
      ctrl->stat_tiles[smid] = __ldg(cstart + group * W + W) - __ldg(cstart + group * W);
      ctrl->stat_cycles[smid] = static_cast<int>(min(static_cast<long long>(0x7fffffff), clock64() - sm.t0));
    }
    // ... and the last CTA of the launch re-arms the claim words of its parity
    __threadfence();
    if (atomicAdd(&ctrl->finished[parity], 1) == G - 1) {
      for (int g = 0; g < G; ++g) ctrl->claimed[parity][g] = 0;
      ctrl->finished[parity] = 0;
      __threadfence();
    }
  }
}

// Warps per CTA: 12 = 3 per SM sub-partition, the most that still leaves each thread the ~168 registers the
// streaming loop needs (a 13th warp would cap one sub-partition at 16384 / 4 / 32 = 128 registers).
constexpr int kSk2Warps = 12;
constexpr int kSk2Depth = 2;  // ring stages per warp of 8 KiB (bf16); fp8 stages are half the bytes, twice the depth

size_t decode_sk2_smem_bytes(bool fp8) {
  return (fp8 ? sizeof(Sk2Smem<kSk2Warps, 2 * kSk2Depth, kKvTileBytesFp8>)
              : sizeof(Sk2Smem<kSk2Warps, kSk2Depth, kKvTileBytesBf16>)) + 1024;
}
int decode_sk2_warps() { return kSk2Warps; }

template <typename K>
static cudaError_t sk2_launch(K kern, int* configured_dev, size_t smem, int threads, const CUtensorMap& tm_k,
                              const CUtensorMap& tm_v, const DecodeParams& p, int grid, bool pdl, cudaStream_t stream) {
  int dev = 0;
  if (cudaError_t e = cudaGetDevice(&dev)) return e;
  if (dev != *configured_dev) {
    if (cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem)))
      return e;
    *configured_dev = dev;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(static_cast<unsigned>(threads));
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, tm_k, tm_v, p);
}

cudaError_t launch_decode_sk2(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const DecodeParams& p, bool fp8,
                              int grid, bool pdl, cudaStream_t stream) {
  static thread_local int dev[2] = {-1, -1};
  if (fp8)
    return sk2_launch(paged_decode_sk2_kernel<kSk2Warps, 2 * kSk2Depth, true>, &dev[1], decode_sk2_smem_bytes(true),
                      kSk2Warps * 32, tm_k, tm_v, p, grid, pdl, stream);
  return sk2_launch(paged_decode_sk2_kernel<kSk2Warps, kSk2Depth, false>, &dev[0], decode_sk2_smem_bytes(false),
                    kSk2Warps * 32, tm_k, tm_v, p, grid, pdl, stream);
}

cudaError_t launch_sk2_reduce(const DecodeParams& p, int grid_decode, bool pdl, cudaStream_t stream) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(p.batch * p.num_kv_heads * p.subgroups));
  cfg.blockDim = dim3(kSk2ReduceWarps * 32);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, sk2_reduce_kernel, p.out, p.out_stride_tok, p.lse, p.batch, p.num_q_heads, p.group,
                            p.subgroups, p.num_kv_heads, grid_decode * kSk2Warps, static_cast<const float*>(p.part_o),
                            static_cast<const float*>(p.part_lse), p.sk2_plan);
}

}  // namespace b200
