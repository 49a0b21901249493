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

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < D; ++s) mbar_init(&sm.bar[warp][s], 1);
    fence_barrier_init();
    if (warp == 0) {
      tma_prefetch_desc(&tm_k);
      tma_prefetch_desc(&tm_v);
    }
  }
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
    if (blockIdx.x == 0) {  // the reduce kernel reads the same plan
      for (int i = threadIdx.x; i <= p.batch; i += W * 32) p.plan_cum[i] = sm.cum[i];
    }
  }

  const int hs_count = p.num_kv_heads * p.subgroups;
  const int64_t total = static_cast<int64_t>(sm.cum[p.batch]) * hs_count;
  const int nw = gridDim.x * W;
  const int tpw = static_cast<int>(max(static_cast<int64_t>(1), (total + nw - 1) / nw));
  const int gw = blockIdx.x * W + warp;
  const int64_t g0 = static_cast<int64_t>(gw) * tpw;
  if (g0 >= total) return;
  const int count = static_cast<int>(min(static_cast<int64_t>(tpw), total - g0));

  // ---- locate the chunk start
  SkPos cs;
  {
    const int key = static_cast<int>(g0 / hs_count);  // cum[b] <= key < cum[b+1]
    int lo_b = 0, hi_b = p.batch;                      // invariant: cum[lo_b] <= key < cum[hi_b]
    while (hi_b - lo_b > 1) {
      const int mid = (lo_b + hi_b) >> 1;
      if (sm.cum[mid] <= key) lo_b = mid; else hi_b = mid;
    }
    cs.b = lo_b;
    cs.n = sm.cum[lo_b + 1] - sm.cum[lo_b];
    const int r = static_cast<int>(g0 - static_cast<int64_t>(sm.cum[lo_b]) * hs_count);
    cs.hs = r / cs.n;
    cs.tile = r % cs.n;
  }
  SkPos pf = cs;

  const SkLane sl = sk_make_lane(lane);
  const uint32_t ring0 = smem_u32(sm.ring[warp][0]);
  // prefetch-front state
  int pf_page0 = (pf.tile >> p.tpp_shift) & ~31;  // block-id window of the item the front is in
  int ids = sk_load_ids(p, pf.b, pf.n, pf_page0, lane);
  int pf_kvh = pf.hs / p.subgroups;
  const uint32_t bar0 = smem_u32(&sm.bar[warp][0]);
  const uint64_t pol = l2_policy_evict_first();  // KV pages are streamed once per step
  int pf_left = count;     // stages the front still has to issue
  int pf_items_ahead = 0;  // item boundaries crossed by the front and not yet by the math (0 or 1)
  bool pf_fresh = false;   // the front stands on tile 0 of an item it has not prepared yet
  int pslot = 0, pending = 0;
  // math state
  int cs_left = count, cslot = 0;
  uint32_t cphase = 0;  // parity of the current pass over the ring
  int seg_start = cs.tile;  // first tile of this warp's segment of the current item
  uint32_t qf[8][2], qf_n[8][2];
  // softmax scale in the log2 domain; fp8: times the K dequantisation scale of the item's kv head
  float scale = p.scale_log2, scale_n = p.scale_log2;
  {
    WorkItem w;
    w.b = cs.b;
    w.kvh = cs.hs / p.subgroups;
    w.sub = cs.hs % p.subgroups;
    if (FP8) {
      sk_load_q_f16(p, w.b, w.kvh, w.sub, lane, qf);
      scale = p.scale_log2 * __ldg(p.k_scale + w.kvh * p.scale_stride);
    } else {
      load_q_frags(p, w, lane, qf);
    }
  }
  WarpState st;
  state_reset(st);
  const int piece_base = gw * 2;

  while (cs_left > 0) {
    // ================= keep the ring full (one issue per iteration in steady state) ==========
    while (pending < D && pf_left > 0) {
      if (pf_fresh) {
        // the front stands on the first tile of an item it has not prepared yet; qf_n can hold one
        // prepared item, so the front never runs two item boundaries ahead of the math
        if (pf_items_ahead == 1) break;
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
        pf_page0 = 0;
        ids = sk_load_ids(p, pf.b, pf.n, 0, lane);
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
      const int piece = piece_base + (seg_start > 0 ? 0 : 1);
      sk_store<FP8>(p, st, cs.b, cs.hs, whole, piece, reinterpret_cast<float*>(sm.ring[warp][cslot]), lane);
      fence_proxy_async_smem();  // the slot was written through the generic proxy; TMA refills it next
      __syncwarp();
      state_reset(st);
      seg_start = 0;
      if (cs_left > 0) {
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
}

// ------------------------------------------------------------------------------------------------
// Merge of the partials left by chunk boundaries. One warp per (sequence, query head) row; rows
// finished by a single warp were already written by it; rows of empty sequences are zeroed here.
// ------------------------------------------------------------------------------------------------
constexpr int kSkReduceRows = 4;

__global__ void __launch_bounds__(kSkReduceRows * 32)
streamk_reduce_kernel(__nv_bfloat16* __restrict__ out, int64_t out_stride_tok, float* __restrict__ lse, int batch,
                      int num_q_heads, int group, int subgroups, int num_kv_heads, int nw,
                      const float* __restrict__ part_o, const float* __restrict__ part_lse,
                      const int32_t* __restrict__ plan_cum, int pdl) {
  if (pdl) griddep_wait();
  const int lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * kSkReduceRows + (threadIdx.x >> 5);
  if (row >= static_cast<int64_t>(batch) * num_q_heads) return;
  const int b = static_cast<int>(row / num_q_heads);
  const int h = static_cast<int>(row % num_q_heads);
  __nv_bfloat16* dst = out + static_cast<int64_t>(b) * out_stride_tok + static_cast<int64_t>(h) * kHeadDim + lane * 4;
  const int c0 = plan_cum[b], n = plan_cum[b + 1] - c0;
  if (n == 0) {  // empty sequence: zero row, LSE = -inf
    *reinterpret_cast<uint2*>(dst) = make_uint2(0u, 0u);
    if (lse != nullptr && lane == 0) lse[row] = -CUDART_INF_F;
    return;
  }
  const int hs_count = num_kv_heads * subgroups;
  const int64_t total = static_cast<int64_t>(plan_cum[batch]) * hs_count;
  const int64_t tpw = max(static_cast<int64_t>(1), (total + nw - 1) / nw);
  const int hig = h % group;
  const int hs = (h / group) * subgroups + hig / kHeadsPerItem;
  const int head = hig % kHeadsPerItem;
  const int64_t t0 = static_cast<int64_t>(c0) * hs_count + static_cast<int64_t>(hs) * n;
  const int w0 = static_cast<int>(t0 / tpw), w1 = static_cast<int>((t0 + n - 1) / tpw);
  if (w0 == w1) return;  // finished by one warp

  auto piece_row = [&](int w) -> int64_t {
    const int slot = 2 * w + ((t0 < static_cast<int64_t>(w) * tpw) ? 0 : 1);
    return static_cast<int64_t>(slot) * kHeadsPerItem + head;
  };
  float mx = -CUDART_INF_F;
  for (int w = w0 + lane; w <= w1; w += 32) mx = fmaxf(mx, part_lse[piece_row(w)]);
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  float denom = 0.f;
  if (mx != -CUDART_INF_F) {
    for (int w = w0; w <= w1; ++w) {
      const int64_t r = piece_row(w);
      const float ls = part_lse[r];
      const float wgt = (ls == -CUDART_INF_F) ? 0.f : fast_exp2((ls - mx) * kLog2e);
      const float4 o = *reinterpret_cast<const float4*>(part_o + r * kHeadDim + lane * 4);
      acc.x = fmaf(o.x, wgt, acc.x);
      acc.y = fmaf(o.y, wgt, acc.y);
      acc.z = fmaf(o.z, wgt, acc.z);
      acc.w = fmaf(o.w, wgt, acc.w);
      denom += wgt;
    }
  }
  const float inv = denom > 0.f ? 1.f / denom : 0.f;
  uint2 pk;
  pk.x = pack_bf16x2(acc.x * inv, acc.y * inv);
  pk.y = pack_bf16x2(acc.z * inv, acc.w * inv);
  *reinterpret_cast<uint2*>(dst) = pk;
  if (lse != nullptr && lane == 0) lse[row] = denom > 0.f ? mx + logf(denom) : -CUDART_INF_F;
}

constexpr int kSkDepthFp8 = 2 * kSkDepth;  // fp8 stages are half the bytes: twice the ring depth

size_t decode_streamk_smem_bytes(bool fp8) {
  return (fp8 ? sizeof(SkSmem<kSkWarps, kSkDepthFp8, kKvTileBytesFp8>)
              : sizeof(SkSmem<kSkWarps, kSkDepth, kKvTileBytesBf16>)) + 1024;
}

template <typename K>
static cudaError_t sk_launch(K kern, int* configured_dev, size_t smem, const CUtensorMap& tm_k, const CUtensorMap& tm_v,
                             const DecodeParams& p, int grid, cudaStream_t stream) {
  int dev = 0;
  if (cudaError_t e = cudaGetDevice(&dev)) return e;
  if (dev != *configured_dev) {
    if (cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem)))
      return e;
    *configured_dev = dev;
  }
  kern<<<grid, kSkWarps * 32, smem, stream>>>(tm_k, tm_v, p);
  return cudaGetLastError();
}

cudaError_t launch_decode_streamk(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const DecodeParams& p, bool fp8,
                                  int grid, cudaStream_t stream) {
  static thread_local int dev_bf16 = -1, dev_fp8 = -1;
  if (fp8)
    return sk_launch(paged_decode_streamk_kernel<kSkWarps, kSkDepthFp8, true>, &dev_fp8,
                     decode_streamk_smem_bytes(true), tm_k, tm_v, p, grid, stream);
  return sk_launch(paged_decode_streamk_kernel<kSkWarps, kSkDepth, false>, &dev_bf16,
                   decode_streamk_smem_bytes(false), tm_k, tm_v, p, grid, stream);
}

cudaError_t launch_streamk_reduce(const DecodeParams& p, int grid_decode, bool pdl, cudaStream_t stream) {
  const int64_t rows = static_cast<int64_t>(p.batch) * p.num_q_heads;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>((rows + kSkReduceRows - 1) / kSkReduceRows));
  cfg.blockDim = dim3(kSkReduceRows * 32);
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
                            p.group, p.subgroups, p.num_kv_heads, nw, static_cast<const float*>(p.part_o),
                            static_cast<const float*>(p.part_lse), static_cast<const int32_t*>(p.plan_cum), pdl_i);
}

}  // namespace b200
