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
//  * No second launch. After its streaming loop the CTA merges the partials its own W consecutive chunks
//    left (sk2_cta_merge: every thread folds one (head, 4 columns) unit with all loads in flight); an item
//    that crosses CTA boundaries is counted per CTA and the LAST CTA to arrive folds the CTA partials.
//    Two short levels whatever the number of pieces (TP4/TP8 shards cut an item into 29-57 pieces, one
//    32k sequence on the whole grid into hundreds), and the cold path runs after the loop, so it cannot
//    raise the register pressure of the loop.
//
// Replaces the decode branch of the reference's stock backends
// (vllm/v1/attention/backends/flashinfer.py:1694-1822, flash_attn.py:758-819) and their split-KV
// combine (vllm_flash_attn flash_fwd_combine; vllm/v1/attention/ops/triton_merge_attn_states.py).
#include "kernels.h"
#include "sk_math.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------
// Plan layout (int32 words): header | desc[nw][4] | ids[nw][32] | cum[batch + 1]
// ------------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ int64_t sk2_plan_desc_off() { return kSk2PlanHdrInts; }
__host__ __device__ __forceinline__ int64_t sk2_plan_ids_off(int nw) { return kSk2PlanHdrInts + 4ll * nw; }
__host__ __device__ __forceinline__ int64_t sk2_plan_cum_off(int nw) { return kSk2PlanHdrInts + 36ll * nw; }

size_t sk2_plan_ints(int batch, int nw) { return static_cast<size_t>(sk2_plan_cum_off(nw)) + batch + 1; }

struct Sk2Pos {
  int b;     // sequence
  int hs;    // kv head * subgroups + head sub-group
  int tile;  // 16-token tile inside the sequence
  int n;     // tiles of sequence b
};

// One CTA. Phase 1: exclusive prefix of tiles per sequence. Phase 2: start position of every
// chunk. Phase 3: the 32 block ids of the page window each chunk starts in.
__global__ void __launch_bounds__(1024)
sk2_plan_kernel(const int32_t* __restrict__ seq_lens, int batch, int hs_count, int nw,
                const int32_t* __restrict__ block_table, int64_t bt_stride, int tiles_per_page, int tpp_shift,
                int32_t* __restrict__ plan) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  int32_t* cum = plan + sk2_plan_cum_off(nw);
  int4* desc = reinterpret_cast<int4*>(plan + sk2_plan_desc_off());
  int32_t* ids = plan + sk2_plan_ids_off(nw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
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
    if (i < batch) cum[i] = carry + wbase + incl - t;
    __syncthreads();
    if (tid == 1023) s_carry = carry + wbase + incl;
    __syncthreads();
  }
  const int tiles = s_carry;
  if (tid == 0) cum[batch] = tiles;
  const int64_t total = static_cast<int64_t>(tiles) * hs_count;
  const int64_t ts64 = (total + nw - 1) / nw;
  const int ts = static_cast<int>(ts64 < 1 ? 1 : ts64);
  if (tid == 0) {
    plan[0] = static_cast<int32_t>(total & 0xffffffffll);
    plan[1] = static_cast<int32_t>(total >> 32);
    plan[2] = ts;
    plan[3] = nw;
    plan[4] = batch;
    plan[5] = hs_count;
    plan[6] = tpp_shift;
    plan[7] = kSk2PlanMagic;
  }
  __syncthreads();  // cum[] is read back below (same CTA: visible after the barrier)
  for (int gw = tid; gw < nw; gw += 1024) {
    const int64_t g0 = static_cast<int64_t>(gw) * ts;
    int4 d = make_int4(batch, 0, 0, 1);
    if (g0 < total) {
      const int key = static_cast<int>(g0 / hs_count);  // cum[b] <= key < cum[b+1]
      int lo = 0, hi = batch;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (cum[mid] <= key) lo = mid; else hi = mid;
      }
      const int n = cum[lo + 1] - cum[lo];
      const int r = static_cast<int>(g0 - static_cast<int64_t>(cum[lo]) * hs_count);
      d = make_int4(lo, r / n, r % n, n);
    }
    desc[gw] = d;
  }
  __syncthreads();
  for (int idx = tid; idx < nw * 32; idx += 1024) {
    const int4 d = desc[idx >> 5];
    const int page0 = (d.z >> tpp_shift) & ~31;
    const int pages = (d.w + tiles_per_page - 1) >> tpp_shift;
    const int pg = page0 + (idx & 31);
    ids[idx] = (d.x < batch && pg < pages) ? block_table[static_cast<int64_t>(d.x) * bt_stride + pg] : 0;
  }
}

cudaError_t launch_sk2_plan(const int32_t* seq_lens, int batch, int hs_count, int nw, const int32_t* block_table,
                            int64_t bt_stride, int tiles_per_page, int tpp_shift, int32_t* plan, cudaStream_t stream) {
  sk2_plan_kernel<<<1, 1024, 0, stream>>>(seq_lens, batch, hs_count, nw, block_table, bt_stride, tiles_per_page,
                                          tpp_shift, plan);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Merge of partials by their LSE (the merge_attn_states formula), by the whole CTA, after its
// streaming loop. A partial is 8 heads x 128 floats + 8 LSEs; thread (head, 4 columns) folds the K
// pieces of one job with every load in flight at once (one L2 round trip per job).
//
// Level 1 — pieces inside this CTA's W consecutive chunks. Job 0: the item the CTA's first chunk
// started inside (pieces: head slots 2c of chunks cfirst ..). Job 1 + w: the item warp w's chunk
// ended inside (pieces: tail slot 2c + 1 of that chunk, then head slots of the chunks after it).
// An item that lies entirely inside the CTA's chunks is finished here; one that crosses a CTA
// boundary leaves a CTA partial in gpart slot 2 * cta + (job 0 ? 0 : 1).
// Level 2 — an item that crosses CTA boundaries is counted once per CTA on cnt2[first CTA]; the CTA
// that arrives last folds the CTA partials (slot 2 * C0 + 1, then 2 * C for the CTAs after it).
// ------------------------------------------------------------------------------------------------
struct Sk2Job {
  int count;       // pieces (0 = no job)
  int first_slot;  // partial slot of piece 0; piece k > 0 lives in slot 2 * (c_start + k)
  int c_start;
  int dst;         // -1 final row, else gpart slot
  int b, hs;
};

template <int KMAX>
__device__ __forceinline__ void sk2_fold(const DecodeParams& p, const float* src_o, const float* src_lse, int count,
                                         int first_slot, int c_start, int dst, int b, int hs, int hh, int col) {
  float m = -CUDART_INF_F, den = 0.f;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int k0 = 0; k0 < count; k0 += KMAX) {
    float l[KMAX];
    float4 o[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      const bool on = k0 + k < count;
      const int64_t row = (static_cast<int64_t>(k0 + k == 0 ? first_slot : 2 * (c_start + k0 + k))) * kHeadsPerItem + hh;
      l[k] = on ? __ldcg(src_lse + row) : -CUDART_INF_F;
      o[k] = on ? __ldcg(reinterpret_cast<const float4*>(src_o + row * kHeadDim + col * 4))
                : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float mn = m;
#pragma unroll
    for (int k = 0; k < KMAX; ++k) mn = fmaxf(mn, l[k]);
    if (mn == -CUDART_INF_F) continue;  // nothing but empty partials so far
    const float a = fast_exp2((m - mn) * kLog2e);  // 0 when m = -inf
    acc.x *= a, acc.y *= a, acc.z *= a, acc.w *= a;
    den *= a;
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      const float w = fast_exp2((l[k] - mn) * kLog2e);  // 0 for -inf
      acc.x = fmaf(o[k].x, w, acc.x);
      acc.y = fmaf(o[k].y, w, acc.y);
      acc.z = fmaf(o[k].z, w, acc.z);
      acc.w = fmaf(o[k].w, w, acc.w);
      den += w;
    }
    m = mn;
  }
  const float inv = den > 0.f ? 1.f / den : 0.f;
  const float lse = den > 0.f ? m + logf(den) : -CUDART_INF_F;
  const float4 v = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
  if (dst >= 0) {
    const int64_t row = static_cast<int64_t>(dst) * kHeadsPerItem + hh;
    *reinterpret_cast<float4*>(p.gpart_o + row * kHeadDim + col * 4) = v;
    if (col == 0) p.gpart_lse[row] = lse;
  } else {
    const int kvh = hs / p.subgroups, sub = hs % p.subgroups;
    const int head_in_group = sub * kHeadsPerItem + hh;
    if (head_in_group < p.group) {
      const int h = kvh * p.group + head_in_group;
      uint2 pk;
      pk.x = pack_bf16x2(v.x, v.y);
      pk.y = pack_bf16x2(v.z, v.w);
      *reinterpret_cast<uint2*>(p.out + static_cast<int64_t>(b) * p.out_stride_tok + static_cast<int64_t>(h) * kHeadDim +
                                col * 4) = pk;
      if (p.lse != nullptr && col == 0) p.lse[static_cast<int64_t>(b) * p.num_q_heads + h] = lse;
    }
  }
}

template <int W>
__device__ __noinline__ void sk2_cta_merge(const DecodeParams& p, const int32_t* plan, int ts, int64_t total,
                                           const int4* s_tail) {
  __shared__ Sk2Job s_job[W + 1];
  __shared__ int s_valid[W + 1];
  __shared__ int s_nvalid;
  __shared__ int s_l2[2][4];  // {first CTA, CTAs, b, hs} of the (at most two) items crossing this CTA's boundaries
  __shared__ int s_l2_last[2];
  const int tid = threadIdx.x;
  const int hs_count = p.num_kv_heads * p.subgroups;
  const int nw = gridDim.x * W;
  const int32_t* cum = plan + sk2_plan_cum_off(nw);
  const int cfirst = blockIdx.x * W, clast = cfirst + W - 1;
  if (tid < 2) {
    s_l2[tid][1] = 0;
    s_l2_last[tid] = 0;
  }
  __syncthreads();  // also: every partial this CTA stored is visible to the whole CTA from here on
  if (tid <= W) {
    Sk2Job j;
    j.count = 0;
    j.first_slot = j.c_start = j.dst = j.b = j.hs = 0;
    int b = -1, hs = 0, n = 1;
    if (tid == 0) {
      const int4 d = __ldg(reinterpret_cast<const int4*>(plan + sk2_plan_desc_off()) + cfirst);
      if (static_cast<int64_t>(cfirst) * ts < total && d.z > 0) b = d.x, hs = d.y, n = d.w;
    } else {
      const int4 t = s_tail[tid - 1];
      if (t.x >= 0) b = t.x, hs = t.y, n = t.z;
    }
    if (b >= 0) {
      const int64_t t0 = static_cast<int64_t>(__ldg(cum + b)) * hs_count + static_cast<int64_t>(hs) * n;
      const int w0 = static_cast<int>(t0 / ts), w1 = static_cast<int>((t0 + n - 1) / ts);
      j.c_start = tid == 0 ? cfirst : cfirst + tid - 1;
      j.count = min(w1, clast) - j.c_start + 1;
      j.first_slot = 2 * j.c_start + (tid == 0 ? 0 : 1);
      j.b = b;
      j.hs = hs;
      const bool final_here = tid > 0 && w1 <= clast;
      j.dst = final_here ? -1 : 2 * static_cast<int>(blockIdx.x) + (tid == 0 ? 0 : 1);
      if (!final_here) {
        const int which = tid == 0 ? 0 : 1;
        s_l2[which][0] = w0 / W;
        s_l2[which][1] = w1 / W - w0 / W + 1;
        s_l2[which][2] = b;
        s_l2[which][3] = hs;
      }
    }
    s_job[tid] = j;
  }
  __syncthreads();
  if (tid == 0) {
    int nv = 0;
    for (int j = 0; j <= W; ++j)
      if (s_job[j].count > 0) s_valid[nv++] = j;
    s_nvalid = nv;
  }
  __syncthreads();
  const int nv = s_nvalid;
  if (nv == 0) return;
  // ---- level 1
  for (int u = tid; u < nv * 256; u += W * 32) {
    const Sk2Job j = s_job[s_valid[u >> 8]];
    sk2_fold<W>(p, p.part_o, p.part_lse, j.count, j.first_slot, j.c_start, j.dst, j.b, j.hs, (u >> 5) & 7, u & 31);
  }
  if (s_l2[0][1] == 0 && s_l2[1][1] == 0) return;  // no item of this CTA crosses a CTA boundary
  // ---- level 2: count this CTA on the crossing items; the last arriver folds the CTA partials
  __threadfence();
  __syncthreads();
  if (tid < 2 && s_l2[tid][1] > 0) {
    const int old = atomicAdd(p.cnt2 + s_l2[tid][0], 1);
    if (old == s_l2[tid][1] - 1) {
      p.cnt2[s_l2[tid][0]] = 0;  // left at zero for the next launch
      s_l2_last[tid] = 1;
    }
  }
  __syncthreads();
#pragma unroll 1
  for (int t = 0; t < 2; ++t) {
    if (!s_l2_last[t]) continue;
    __threadfence();
    const int c0 = s_l2[t][0];
    if (tid < 256)
      sk2_fold<8>(p, p.gpart_o, p.gpart_lse, s_l2[t][1], 2 * c0 + 1, c0, -1, s_l2[t][2], s_l2[t][3], (tid >> 5) & 7,
                  tid & 31);
  }
}

// Rows of empty sequences (seq_len <= 0, e.g. CUDA-graph padding): zero output, LSE = -inf.
__device__ __noinline__ void sk2_zero_rows(const DecodeParams& p, int b, int lane) {
  for (int h = 0; h < p.num_q_heads; ++h) {
    *reinterpret_cast<uint2*>(p.out + static_cast<int64_t>(b) * p.out_stride_tok + static_cast<int64_t>(h) * kHeadDim +
                              lane * 4) = make_uint2(0u, 0u);
    if (p.lse != nullptr && lane == 0) p.lse[static_cast<int64_t>(b) * p.num_q_heads + h] = -CUDART_INF_F;
  }
}

template <int W, int D, int SB>
struct Sk2Smem {
  alignas(1024) uint8_t ring[W][D][SB];  // SB = bytes of one stage (K tile + V tile)
  alignas(8) uint64_t bar[W][D];         // "stage landed" mbarriers, one per ring slot
  alignas(16) int4 tail[W];              // per warp: the item its chunk ended inside {b, hs, n} (b < 0: none)
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
  const int gw = blockIdx.x * W + warp;
  if (p.trace != nullptr && lane == 0) p.trace[gw * 4 + 0] = sk2_globaltimer();  // debug: entry stamp
  bool waited = false;
  if (p.flags & kSk2FlagWaitFirst) {  // the plan was written by the kernel right before this one
    griddep_wait();
    waited = true;
  }
  // ---- the warp's work: one round of independent loads
  const int32_t* plan = p.sk2_plan;
  const int4 hdr = __ldg(reinterpret_cast<const int4*>(plan));
  const int4 hdr2 = __ldg(reinterpret_cast<const int4*>(plan) + 1);
  const int4 d4 = __ldg(reinterpret_cast<const int4*>(plan + sk2_plan_desc_off()) + gw);
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
  __syncwarp();
  // the next kernel of the chain may start its own prologue (it cannot get an SM before this CTA exits)
  griddep_launch_dependents();

  const int hs_count = p.num_kv_heads * p.subgroups;
  const int64_t total = (static_cast<int64_t>(hdr.y) << 32) | static_cast<uint32_t>(hdr.x);
  const int ts = hdr.z;
  if (hdr.w != nw || hdr2.x != p.batch || hdr2.y != hs_count || hdr2.z != p.tpp_shift || hdr2.w != kSk2PlanMagic)
    __trap();  // the plan in the workspace was made for another launch shape: fail loudly
  const int64_t g0 = static_cast<int64_t>(gw) * ts;
  const int cnt = static_cast<int>(max(static_cast<int64_t>(0), min(static_cast<int64_t>(ts), total - g0)));
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
  int tail_b = -1, tail_hs = 0, tail_n = 0;  // the item this chunk ends inside, if it leaves a partial there
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
        griddep_wait();
        waited = true;
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
      griddep_wait();
      waited = true;
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
    if (p.trace != nullptr && cs_left == cnt && lane == 0) p.trace[gw * 4 + 1] = sk2_globaltimer();  // first stage landed
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
      if (!whole && seg_start == 0) {  // cut at the chunk end (a segment with seg_start > 0 is the head piece)
        tail_b = cs.b;
        tail_hs = cs.hs;
        tail_n = cs.n;
      }
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
  if (!waited) griddep_wait();
  for (int b = gw; b < p.batch; b += nw)
    if (__ldg(p.seq_lens + b) <= 0) sk2_zero_rows(p, b, lane);
  // ---- merge of the partials (sk2_cta_merge): after the streaming loop, so none of the loop's state is live
  // across it; nothing is lost by merging late: every other piece of an item is finished at the end of its
  // own warp's chunk as well.
  if (lane == 0) sm.tail[warp] = make_int4(tail_b, tail_hs, tail_n, 0);
  if (p.trace != nullptr && lane == 0) {  // debug: end of this warp's streaming
    const uint64_t t_end = sk2_globaltimer();
    if (cnt == 0) p.trace[gw * 4 + 1] = t_end;
    p.trace[gw * 4 + 2] = t_end;
  }
  sk2_cta_merge<W>(p, plan, ts, total, sm.tail);
  if (p.trace != nullptr && lane == 0) p.trace[gw * 4 + 3] = sk2_globaltimer();  // debug: merges done
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

}  // namespace b200
