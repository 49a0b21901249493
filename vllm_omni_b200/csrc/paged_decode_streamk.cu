// Paged decode attention for sm_100a, bf16 KV cache — "stream-K" variant (the default).
//
// All 16-token KV tiles of the step form one global stream, ordered (sequence, kv head [, head
// sub-group], tile). The stream is cut into NW = SMs x W equal contiguous chunks, one per WARP of
// a persistent grid (one CTA per SM): perfect balance for any batch / raggedness without a split
// heuristic, a host-independent grid (CUDA-graph safe) and no cross-warp synchronisation at all.
// Each warp streams its chunk through a private ring of D 8-KiB stages (K tile + V tile). The
// warp fills its own ring: its lane 0 issues two TMA tile loads per stage (5-D tensor maps over
// the cache: [64 d_lo][slot][2 d_hi][head][block], box 64 x 16 x 2, SWIZZLE_128B — any cache
// strides) completing on a per-slot mbarrier, so loading costs three instructions per 8 KiB, no
// LSU traffic and no producer/consumer hand-off (self-issued TMA reaches ~7.1 TB/s on random
// 8 KiB pages in scripts/probe/stream_probe.cu; a producer-warp ring only ~3.4 TB/s). The warp
// then runs the tensor-core math on the swizzled tiles. The prefetch front walks the same stream
// D stages ahead of the math, straight across work-item boundaries.
//
// A (sequence, kv head) item that lies inside one chunk is finished by that warp (bf16 output +
// LSE written directly). An item cut by chunk boundaries leaves one normalised partial (O, LSE)
// per warp that touched it — at most two partials per warp, so the workspace is a fixed
// 2 x NW x 8 x (128 + 1) floats — and streamk_reduce_kernel merges them by their LSE.
//
// Replaces the decode branch of the reference's stock backends
// (vllm/v1/attention/backends/flashinfer.py:1694-1822, flash_attn.py:758-819) and their split-KV
// combine (vllm_flash_attn flash_fwd_combine; vllm/v1/attention/ops/triton_merge_attn_states.py).
#include "kernels.h"
#include "sk_math.cuh"

namespace b200 {

template <int W, int D, int SB>
struct SkSmem {
  alignas(1024) uint8_t ring[W][D][SB];  // SB = bytes of one stage (K tile + V tile)
  int32_t cum[kSkMaxBatch + 1];  // exclusive prefix of tiles per sequence
  int32_t slen[kSkMaxBatch];     // clamped sequence lengths
  int32_t warp_sums[32];
  alignas(8) uint64_t bar[W][D];  // "stage landed" mbarriers, one per ring slot
};

// Position in the tile stream.
struct SkPos {
  int b;     // sequence
  int hs;    // kv head * subgroups + head sub-group
  int tile;  // 16-token tile inside the sequence
  int n;     // tiles of sequence b
};

__device__ __forceinline__ void sk_advance(const int32_t* cum, int batch, int hs_count, SkPos& it) {
  if (++it.tile < it.n) return;
  it.tile = 0;
  if (++it.hs < hs_count) return;
  it.hs = 0;
  do {
    ++it.b;
  } while (it.b < batch && cum[it.b + 1] == cum[it.b]);
  it.n = it.b < batch ? cum[it.b + 1] - cum[it.b] : 1;
}

// 32 consecutive block ids of sequence b starting at page `page0`, one per lane.
__device__ __forceinline__ int sk_load_ids(const DecodeParams& p, int b, int n_tiles, int page0, int lane) {
  const int pages = (n_tiles + p.tiles_per_page - 1) >> p.tpp_shift;
  const int pg = page0 + lane;
  return (b < p.batch && pg < pages) ? __ldg(p.block_table + static_cast<int64_t>(b) * p.bt_stride + pg) : 0;
}

// Locate stream position g (in tiles x head-groups, see the file header) -> (sequence, hs, tile).
__device__ __forceinline__ SkPos sk_locate(const int32_t* cum, int batch, int hs_count, int64_t g) {
  SkPos o;
  const int key = static_cast<int>(g / hs_count);  // cum[b] <= key < cum[b+1]
  int lo_b = 0, hi_b = batch;                      // invariant: cum[lo_b] <= key < cum[hi_b]
  while (hi_b - lo_b > 1) {
    const int mid = (lo_b + hi_b) >> 1;
    if (cum[mid] <= key) lo_b = mid; else hi_b = mid;
  }
  o.b = lo_b;
  o.n = cum[lo_b + 1] - cum[lo_b];
  const int r = static_cast<int>(g - static_cast<int64_t>(cum[lo_b]) * hs_count);
  o.hs = r / o.n;
  o.tile = r % o.n;
  return o;
}

__device__ __forceinline__ uint64_t sk_globaltimer() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

template <int W, int D, bool FP8>
__global__ void __launch_bounds__(W * 32, 1)
paged_decode_streamk_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                            const DecodeParams p) {
  constexpr int SB = FP8 ? kKvTileBytesFp8 : kKvTileBytesBf16;  // stage bytes
  constexpr int VOFF = SB / 2;                                   // V tile behind the K tile
  extern __shared__ uint8_t smem_raw[];
  SkSmem<W, D, SB>& sm =
      *reinterpret_cast<SkSmem<W, D, SB>*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint64_t t_entry = p.trace != nullptr ? sk_globaltimer() : 0;

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < D; ++s) mbar_init(&sm.bar[warp][s], 1);
    fence_barrier_init();
    if (warp == 0) {
      tma_prefetch_desc(&tm_k);
      tma_prefetch_desc(&tm_v);
    }
  }
  // Programmatic dependent launch. Only this library's own kernels release their dependents early
  // (cache write, decode, reduce) and none of them writes seq_lens, the block table or q, so the
  // plan below may overlap the previous kernel; griddepcontrol.wait comes right before the first
  // access to what they DO write — the KV cache, the workspace plan and the output.
  griddep_launch_dependents();
  // ---- per-CTA plan: tiles per sequence and their exclusive prefix (batch <= kSkMaxBatch)
  {
    const int per = (p.batch + W * 32 - 1) / (W * 32);
    const int i0 = threadIdx.x * per;
    int local = 0;
    for (int i = i0; i < min(p.batch, i0 + per); ++i) {
      const int s = max(0, __ldg(p.seq_lens + i));
      sm.slen[i] = s;
      local += (s + kTileTokens - 1) / kTileTokens;
    }
    int incl = local;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += v;
    }
    if (lane == 31) sm.warp_sums[warp] = incl;
    __syncthreads();
    int base = 0;
    for (int ww = 0; ww < warp; ++ww) base += sm.warp_sums[ww];
    int run = base + incl - local;
    for (int i = i0; i < min(p.batch, i0 + per); ++i) {
      sm.cum[i] = run;
      run += (sm.slen[i] + kTileTokens - 1) / kTileTokens;
    }
    if (threadIdx.x == W * 32 - 1) sm.cum[p.batch] = base + incl;
    __syncthreads();
    if (blockIdx.x == 0) {  // the reduce kernel reads the same plan (a previous reduce may still be reading it)
      griddep_wait();
      for (int i = threadIdx.x; i <= p.batch; i += W * 32) p.plan_cum[i] = sm.cum[i];
    }
  }

  const int hs_count = p.num_kv_heads * p.subgroups;
  const int nw = gridDim.x * W;
  const SkSchedule sc = sk_schedule(static_cast<int64_t>(sm.cum[p.batch]) * hs_count, nw, p.sk_dyn_shift);
  const int gw = blockIdx.x * W + warp;

  const SkLane sl = sk_make_lane(lane);
  const uint32_t ring0 = smem_u32(sm.ring[warp][0]);
  const uint32_t bar0 = smem_u32(&sm.bar[warp][0]);
  const uint64_t pol = l2_policy_evict_first();  // KV pages are streamed once per step

  // ---- chunk bookkeeping. The warp's first chunk is its static one; further chunks are claimed
  // from the global counter. `nxt_*` is the one chunk the prefetch front may be ahead of the math.
  bool front_ahead = false;  // the front stands in chunk nxt_*, the math still in the previous one
  bool pf_done = false;      // no chunk left to claim
  bool first = true;         // the static chunk has not been handed out yet
  bool waited = false;       // griddepcontrol.wait executed
  SkPos nxt_pos{0, 0, 0, 1};
  int nxt_count = 0, nxt_id = 0;
  int claim_val = 0;
  bool claim_out = false;  // an atomic claim is in flight (result in lane 0's claim_val)

  // prefetch-front state
  SkPos pf{0, 0, 0, 1};
  int pf_left = 0;         // tiles of the front's chunk still to issue
  int pf_page0 = 0, ids = 0, pf_kvh = 0;
  int pf_items_ahead = 0;  // item/chunk starts prepared by the front and not yet taken by the math (0 or 1)
  bool pf_fresh = false;   // the front stands on a tile whose item it has not prepared yet
  int pslot = 0, pending = 0;
  // math state
  SkPos cs{0, 0, 0, 1};
  int cs_left = 0, cslot = 0, cur_id = 0;
  uint32_t cphase = 0;  // parity of the current pass over the ring
  int seg_start = 0;    // first tile of this warp's segment of the current item
  uint32_t qf[8][2], qf_n[8][2];
#pragma unroll
  for (int kk = 0; kk < 8; ++kk) qf[kk][0] = qf[kk][1] = qf_n[kk][0] = qf_n[kk][1] = 0u;
  // softmax scale in the log2 domain; fp8: times the K dequantisation scale of the item's kv head
  float scale = p.scale_log2, scale_n = p.scale_log2;
  WarpState st;
  state_reset(st);
  uint64_t t_static_end = 0;

  while (true) {
    // ================= keep the ring full (one issue per iteration in steady state) ==========
    while (pending < D) {
      // claim the next dynamic chunk shortly before the front runs out of its current one, so the
      // atomic's round trip hides behind the tiles still to be issued
      if (sc.dyn && !first && !claim_out && !pf_done && pf_left <= kSkClaimLead) {
        if (lane == 0) claim_val = atomicAdd(p.sk_ctrl, 1);
        claim_out = true;
      }
      if (pf_left == 0) {
        if (pf_done || front_ahead) break;  // one chunk descriptor can be outstanding
        int64_t g0;
        int cnt, id;
        if (first) {
          first = false;
          g0 = static_cast<int64_t>(gw) * sc.ts;
          cnt = static_cast<int>(max(static_cast<int64_t>(0), min(static_cast<int64_t>(sc.ts), sc.S - g0)));
          id = gw;
          if (cnt == 0) {
            if (!sc.dyn) pf_done = true;
            continue;  // nothing static for this warp: go claim (or stop)
          }
        } else {
          if (!sc.dyn) {
            pf_done = true;
            break;
          }
          const int j = __shfl_sync(0xffffffffu, claim_val, 0);
          claim_out = false;
          if (j >= sc.nd) {
            pf_done = true;
            break;
          }
          g0 = sc.S + static_cast<int64_t>(j) * sc.cd;
          cnt = static_cast<int>(min(static_cast<int64_t>(sc.cd), sc.total - g0));
          id = nw + j;
        }
        pf = sk_locate(sm.cum, p.batch, hs_count, g0);
        pf_left = cnt;
        nxt_pos = pf;
        nxt_count = cnt;
        nxt_id = id;
        front_ahead = true;
        pf_fresh = true;
      }
      if (pf_fresh) {
        // the front stands on the first tile of an item (or chunk) it has not prepared yet; qf_n can
        // hold one prepared item, so the front never runs two starts ahead of the math
        if (pf_items_ahead == 1) break;
        if (!waited) {  // first touch of q / the KV cache: the previous kernel must be complete
          griddep_wait();
          waited = true;
        }
        WorkItem w;
        w.b = pf.b;
        w.kvh = pf.hs / p.subgroups;
        w.sub = pf.hs % p.subgroups;
        if (FP8) {
          sk_load_q_f16(p, w.b, w.kvh, w.sub, lane, qf_n);
          scale_n = p.scale_log2 * __ldg(p.k_scale + w.kvh * p.scale_stride);
        } else {
          load_q_frags(p, w, lane, qf_n);
        }
        pf_page0 = (pf.tile >> p.tpp_shift) & ~31;
        ids = sk_load_ids(p, pf.b, pf.n, pf_page0, lane);
        pf_kvh = w.kvh;
        pf_items_ahead = 1;
        pf_fresh = false;
      }
      const int page = pf.tile >> p.tpp_shift;
      if (page >= pf_page0 + 32) {
        pf_page0 = page & ~31;
        ids = sk_load_ids(p, pf.b, pf.n, pf_page0, lane);
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
      sk_advance(sm.cum, p.batch, hs_count, pf);
      pf_fresh = (pf.tile == 0);
    }

    if (cs_left == 0) {
      // the math finished its chunk (the segment was stored with its last tile): take the next one
      if (!front_ahead) break;  // nothing left
      if (p.trace != nullptr && cur_id < nw && t_static_end == 0 && nxt_id >= nw) t_static_end = sk_globaltimer();
      cs = nxt_pos;
      cs_left = nxt_count;
      cur_id = nxt_id;
      seg_start = cs.tile;
      front_ahead = false;
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        qf[kk][0] = qf_n[kk][0];
        qf[kk][1] = qf_n[kk][1];
      }
      scale = scale_n;
      pf_items_ahead = 0;
      continue;
    }

    // ================= math on the oldest stage =================
    mbar_wait(&sm.bar[warp][cslot], cphase);
    const bool item_done = (cs.tile + 1 == cs.n);
    {
      const int valid = item_done ? sm.slen[cs.b] - cs.tile * kTileTokens : kTileTokens;
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
      const int piece = cur_id * 2 + (seg_start > 0 ? 0 : 1);
      sk_store<FP8>(p, st, cs.b, cs.hs, whole, piece, reinterpret_cast<float*>(sm.ring[warp][cslot]), lane);
      fence_proxy_async_smem();  // the slot was written through the generic proxy; TMA refills it next
      __syncwarp();
      state_reset(st);
      seg_start = 0;
      if (cs_left > 0) {  // next item of the same chunk: the front prepared it already
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          qf[kk][0] = qf_n[kk][0];
          qf[kk][1] = qf_n[kk][1];
        }
        scale = scale_n;
        pf_items_ahead = 0;
      }
    }
    if (++cslot == D) {
      cslot = 0;
      cphase ^= 1u;
    }
    sk_advance(sm.cum, p.batch, hs_count, cs);
  }

  // ---- the last warp to finish re-arms the claim counter for the next launch
  if (sc.dyn && lane == 0) {
    __threadfence();
    const int old = atomicAdd(p.sk_ctrl + 1, 1);
    if (old == nw - 1) {
      p.sk_ctrl[0] = 0;
      p.sk_ctrl[1] = 0;
      __threadfence();
    }
  }
  if (p.trace != nullptr && lane == 0) {
    const uint64_t t_end = sk_globaltimer();
    p.trace[gw * 3 + 0] = t_entry;
    p.trace[gw * 3 + 1] = t_static_end != 0 ? t_static_end : t_end;
    p.trace[gw * 3 + 2] = t_end;
  }
}

// ------------------------------------------------------------------------------------------------
// Merge of the partials left by chunk boundaries (the separate log-sum-exp reduce kernel). One CTA
// of 8 warps per (sequence, kv head, head sub-group) item: a partial is one contiguous 4 KiB block
// (8 heads x 128 floats) plus 8 LSEs, so warp w merges the pieces k = w, w + 8, ... for all 8 heads
// with 8 independent 512 B loads in flight per piece, and the 8 warps are folded through shared
// memory. Items finished by a single warp were already written by it; rows of empty sequences are
// zeroed here.
// ------------------------------------------------------------------------------------------------
constexpr int kSkReduceWarps = 8;

__global__ void __launch_bounds__(kSkReduceWarps * 32)
streamk_reduce_kernel(__nv_bfloat16* __restrict__ out, int64_t out_stride_tok, float* __restrict__ lse, int batch,
                      int num_q_heads, int group, int subgroups, int num_kv_heads, int nw, int dyn_shift,
                      const float* __restrict__ part_o, const float* __restrict__ part_lse,
                      const int32_t* __restrict__ plan_cum, int pdl) {
  __shared__ float s_acc[kSkReduceWarps][kHeadsPerItem][kHeadDim];
  __shared__ float s_max[kSkReduceWarps][kHeadsPerItem];
  __shared__ float s_den[kSkReduceWarps][kHeadsPerItem];
  if (pdl) {
    griddep_wait();
    griddep_launch_dependents();
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int hs_count = num_kv_heads * subgroups;
  const int b = blockIdx.x / hs_count, hs = blockIdx.x % hs_count;
  const int kvh = hs / subgroups, sub = hs % subgroups;
  const int heads_here = min(kHeadsPerItem, group - sub * kHeadsPerItem);
  const int h_first = kvh * group + sub * kHeadsPerItem;  // first query head of the item
  const int c0 = plan_cum[b], n = plan_cum[b + 1] - c0;
  if (n == 0) {  // empty sequence: zero rows, LSE = -inf
    if (warp < heads_here) {
      const int h = h_first + warp;
      *reinterpret_cast<uint2*>(out + static_cast<int64_t>(b) * out_stride_tok + static_cast<int64_t>(h) * kHeadDim +
                                lane * 4) = make_uint2(0u, 0u);
      if (lse != nullptr && lane == 0) lse[static_cast<int64_t>(b) * num_q_heads + h] = -CUDART_INF_F;
    }
    return;
  }
  const SkSchedule sc = sk_schedule(static_cast<int64_t>(plan_cum[batch]) * hs_count, nw, dyn_shift);
  // the item's tiles [t0, t1) in stream order and the chunks that overlap them
  const int64_t t0 = static_cast<int64_t>(c0) * hs_count + static_cast<int64_t>(hs) * n, t1 = t0 + n;
  int w0 = 0, ns = 0;  // static chunks w0 .. w0 + ns - 1
  if (t0 < sc.S) {
    w0 = static_cast<int>(t0 / sc.ts);
    ns = static_cast<int>(((t1 < sc.S ? t1 : sc.S) - 1) / sc.ts) - w0 + 1;
  }
  int j0 = 0, nj = 0;  // dynamic chunks j0 .. j0 + nj - 1
  if (t1 > sc.S) {
    j0 = static_cast<int>(((t0 > sc.S ? t0 : sc.S) - sc.S) / sc.cd);
    nj = static_cast<int>((t1 - 1 - sc.S) / sc.cd) - j0 + 1;
  }
  const int pieces = ns + nj;
  if (pieces == 1) return;  // finished by one warp

  auto piece_slot = [&](int k) -> int64_t {  // k-th overlapping chunk -> its partial slot
    int cid;
    int64_t cstart;
    if (k < ns) {
      cid = w0 + k;
      cstart = static_cast<int64_t>(cid) * sc.ts;
    } else {
      const int j = j0 + (k - ns);
      cid = nw + j;
      cstart = sc.S + static_cast<int64_t>(j) * sc.cd;
    }
    return static_cast<int64_t>(2 * cid + (t0 < cstart ? 0 : 1));
  };

  // ---- one pass: warp w folds the pieces k = w, w + 8, ... of all heads with a running reference
  // (LSE and O of a piece are loaded together: one L2 round trip per piece instead of two)
  float4 acc[kHeadsPerItem];
  float den[kHeadsPerItem], m[kHeadsPerItem];
#pragma unroll
  for (int hh = 0; hh < kHeadsPerItem; ++hh) {
    acc[hh] = make_float4(0.f, 0.f, 0.f, 0.f);
    den[hh] = 0.f;
    m[hh] = -CUDART_INF_F;
  }
  for (int k = warp; k < pieces; k += kSkReduceWarps) {
    const int64_t r0 = piece_slot(k) * kHeadsPerItem;
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
    for (int w = 1; w < kSkReduceWarps; ++w) mfin = fmaxf(mfin, s_max[w][warp]);
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    float dsum = 0.f;
#pragma unroll
    for (int w = 0; w < kSkReduceWarps; ++w) {
      const float mw = s_max[w][warp];
      const float sc_w = (mw == -CUDART_INF_F) ? 0.f : fast_exp2((mw - mfin) * kLog2e);
      const float4 v = *reinterpret_cast<const float4*>(&s_acc[w][warp][lane * 4]);
      a.x = fmaf(v.x, sc_w, a.x);
      a.y = fmaf(v.y, sc_w, a.y);
      a.z = fmaf(v.z, sc_w, a.z);
      a.w = fmaf(v.w, sc_w, a.w);
      dsum = fmaf(s_den[w][warp], sc_w, dsum);
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

constexpr int kSkDepthFp8 = 2 * kSkDepth;  // fp8 stages are half the bytes: twice the ring depth

size_t decode_streamk_smem_bytes(bool fp8) {
  return (fp8 ? sizeof(SkSmem<kSkWarps, kSkDepthFp8, kKvTileBytesFp8>)
              : sizeof(SkSmem<kSkWarps, kSkDepth, kKvTileBytesBf16>)) + 1024;
}

template <typename K>
static cudaError_t sk_launch(K kern, int* configured_dev, size_t smem, const CUtensorMap& tm_k, const CUtensorMap& tm_v,
                             const DecodeParams& p, int grid, bool pdl, cudaStream_t stream) {
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
  cfg.blockDim = dim3(kSkWarps * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, tm_k, tm_v, p);
}

cudaError_t launch_decode_streamk(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const DecodeParams& p, bool fp8,
                                  int grid, bool pdl, cudaStream_t stream) {
  static thread_local int dev_bf16 = -1, dev_fp8 = -1;
  if (fp8)
    return sk_launch(paged_decode_streamk_kernel<kSkWarps, kSkDepthFp8, true>, &dev_fp8,
                     decode_streamk_smem_bytes(true), tm_k, tm_v, p, grid, pdl, stream);
  return sk_launch(paged_decode_streamk_kernel<kSkWarps, kSkDepth, false>, &dev_bf16,
                   decode_streamk_smem_bytes(false), tm_k, tm_v, p, grid, pdl, stream);
}

cudaError_t launch_streamk_reduce(const DecodeParams& p, int grid_decode, bool pdl, cudaStream_t stream) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(p.batch * p.num_kv_heads * p.subgroups));
  cfg.blockDim = dim3(kSkReduceWarps * 32);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  const int pdl_i = pdl ? 1 : 0;
  const int nw = grid_decode * kSkWarps;
  return cudaLaunchKernelEx(&cfg, streamk_reduce_kernel, p.out, p.out_stride_tok, p.lse, p.batch, p.num_q_heads,
                            p.group, p.subgroups, p.num_kv_heads, nw, p.sk_dyn_shift,
                            static_cast<const float*>(p.part_o), static_cast<const float*>(p.part_lse),
                            static_cast<const int32_t*>(p.plan_cum), pdl_i);
}

}  // namespace b200
