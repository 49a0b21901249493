// Split-K paged decode attention for sm_100a, bf16 KV cache.
//
//  * paged_decode_tma_kernel    — persistent (one CTA per SM), warp-specialised: one producer
//    warp walks the block table and issues TMA tile loads (cp.async.bulk.tensor, SWIZZLE_128B)
//    into an mbarrier ring; W consumer warps each take whole stages (32 tokens) round-robin,
//    run the MMA math with a private online softmax and are merged through shared memory at the
//    end of a work item. The ring decouples HBM streaming from the per-item epilogues.
//  * paged_decode_cpasync_kernel — one CTA per work item, each warp stages its own pages with
//    16-byte cp.async into a private double buffer (same swizzled layout, same math).
//  * lse_reduce_kernel           — merges the per-split partial outputs by their LSE.
//
// Replaces the decode branch of the reference's stock backends
// (vllm/v1/attention/backends/flashinfer.py:1694-1822, flash_attn.py:758-819).
#include "kernels.h"

namespace b200 {

constexpr int kStageTiles = 2;                                   // 32 tokens per stage
constexpr int kStageBytes = kStageTiles * kKvTileBytesBf16;      // 16 KiB
constexpr int kTmaConsumerWarps = 8;
constexpr int kTmaStages = 10;                                    // 160 KiB ring
constexpr int kCpWarps = 4;

// Opt in to > 48 KiB of dynamic shared memory once per (thread, device).
template <typename K>
static cudaError_t ensure_dynamic_smem(K kern, size_t bytes, int* configured_dev) {
  int dev = 0;
  if (cudaError_t e = cudaGetDevice(&dev)) return e;
  if (dev == *configured_dev) return cudaSuccess;
  if (cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           static_cast<int>(bytes)))
    return e;
  *configured_dev = dev;
  return cudaSuccess;
}

// ------------------------------------------------------------------------------------------------
// TMA variant
// ------------------------------------------------------------------------------------------------
template <int W, int NS>
struct TmaSmem {
  alignas(1024) uint8_t stages[NS][kStageBytes];
  alignas(16) MergeSmem<W> merge;
  alignas(8) uint64_t full[NS];
  alignas(8) uint64_t empty[NS];
};

__device__ __forceinline__ int pick3(int sel, int a, int b, int c) {
  return sel == 0 ? a : (sel == 1 ? b : c);
}

template <int W, int NS>
__global__ void __launch_bounds__((W + 1) * 32, 1)
paged_decode_tma_kernel(const __grid_constant__ CUtensorMap tm_k,
                        const __grid_constant__ CUtensorMap tm_v, const CoordPerm perm,
                        const DecodeParams p) {
  extern __shared__ uint8_t smem_raw[];
  // keep the pointer derived from the __shared__ array so the compiler emits LDS/STS
  TmaSmem<W, NS>& sm =
      *reinterpret_cast<TmaSmem<W, NS>*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], 1);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
  }
  __syncthreads();

  if (warp == 0) {
    // ============================ producer warp ============================
    const uint64_t pol = l2_policy_evict_first();
    const int tpp = p.tiles_per_page;
    // lane -> TMA op of a stage: tile (0/1), K or V, d half (0/1)
    const int op_tile = lane >> 2, op_v = (lane >> 1) & 1, op_half = lane & 1;
    uint32_t n = 0;  // stage counter over the whole kernel
    for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
      const WorkItem w = decode_item(p, item);
      const int ntiles = w.tile_end - w.tile_begin;
      if (ntiles <= 0) continue;
      const int nst = (ntiles + kStageTiles - 1) / kStageTiles;
      const int32_t* bt = p.block_table + static_cast<int64_t>(w.b) * p.bt_stride;
      const int pages_total = (w.seq_len + p.block_size - 1) / p.block_size;
      // 2 x 32-page window of block ids held across the warp
      int pb = w.tile_begin / tpp;
      int ids = (pb + lane < pages_total) ? __ldg(bt + pb + lane) : 0;
      int ids_next = (pb + 32 + lane < pages_total) ? __ldg(bt + pb + 32 + lane) : 0;
      for (int i = 0; i < nst; ++i, ++n) {
        const int slot = n % NS;
        const uint32_t par = ((n / NS) & 1) ^ 1;
        const int tile0 = w.tile_begin + i * kStageTiles;
        const int nt = min(kStageTiles, w.tile_end - tile0);
        const int tile = tile0 + op_tile;
        const int rel = tile / tpp - pb;  // < 64 by construction
        const int blk_a = __shfl_sync(0xffffffffu, ids, rel & 31);
        const int blk_b = __shfl_sync(0xffffffffu, ids_next, rel & 31);
        const int blk = rel < 32 ? blk_a : blk_b;
        if (lane == 0) {
          mbar_wait(&sm.empty[slot], par);
          mbar_arrive_expect_tx(&sm.full[slot], static_cast<uint32_t>(nt) * kKvTileBytesBf16);
        }
        __syncwarp();
        if (lane < nt * 4) {
          uint8_t* dst = sm.stages[slot] + op_tile * kKvTileBytesBf16 + op_v * kTileBytesBf16 +
                         op_half * kHalfBytes;
          const int slot0 = (tile % tpp) * kTileTokens;
          const int c1 = pick3(perm.dim1, slot0, w.kvh, blk);
          const int c2 = pick3(perm.dim2, slot0, w.kvh, blk);
          const int c3 = pick3(perm.dim3, slot0, w.kvh, blk);
          tma_load_4d(dst, op_v ? &tm_v : &tm_k, &sm.full[slot], op_half * 64, c1, c2, c3, pol);
        }
        if ((tile0 + kStageTiles) / tpp >= pb + 32) {  // next stage starts in the next window
          pb += 32;
          ids = ids_next;
          ids_next = (pb + 32 + lane < pages_total) ? __ldg(bt + pb + 32 + lane) : 0;
        }
      }
    }
  } else {
    // ============================ consumer warps ============================
    const int cw = warp - 1;
    const LaneOffsets lo = make_lane_offsets(lane);
    uint32_t n_base = 0;
    WarpState st;
    for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
      const WorkItem w = decode_item(p, item);
      if (!item_active(w)) continue;
      const int ntiles = w.tile_end - w.tile_begin;
      const int nst = (ntiles + kStageTiles - 1) / kStageTiles;
      uint32_t qf[8][2];
      load_q_frags(p, w, lane, qf);
      state_reset(st);
      const int i0 = (cw + W - static_cast<int>(n_base % W)) % W;
      for (int i = i0; i < nst; i += W) {
        const uint32_t n = n_base + i;
        const int slot = n % NS;
        const uint32_t par = (n / NS) & 1;
        const int tile0 = w.tile_begin + i * kStageTiles;
        const int nt = min(kStageTiles, w.tile_end - tile0);
        const int valid = min(nt * kTileTokens, w.seq_len - tile0 * kTileTokens);
        const bool need_mask = valid < nt * kTileTokens;
        mbar_wait(&sm.full[slot], par);
        if (nt == 2)
          consume_stage_bf16<2>(sm.stages[slot], valid, need_mask, qf, lo, p.scale_log2, st, lane);
        else
          consume_stage_bf16<1>(sm.stages[slot], valid, need_mask, qf, lo, p.scale_log2, st, lane);
        if (need_mask) fence_proxy_async_smem();  // zeroed V rows vs. the next TMA write
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[slot]);
      }
      n_base += nst;
      merge_and_store<W>(p, w, st, sm.merge, cw, lane, 1);
    }
  }
}

size_t decode_tma_smem_bytes() { return sizeof(TmaSmem<kTmaConsumerWarps, kTmaStages>) + 1024; }

cudaError_t launch_decode_tma(const CUtensorMap& tm_k, const CUtensorMap& tm_v, CoordPerm perm,
                              const DecodeParams& p, int grid, cudaStream_t stream) {
  auto kern = paged_decode_tma_kernel<kTmaConsumerWarps, kTmaStages>;
  static thread_local int configured_dev = -1;
  const size_t smem = decode_tma_smem_bytes();
  if (cudaError_t e = ensure_dynamic_smem(kern, smem, &configured_dev)) return e;
  kern<<<grid, (kTmaConsumerWarps + 1) * 32, smem, stream>>>(tm_k, tm_v, perm, p);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// cp.async variant
// ------------------------------------------------------------------------------------------------
template <int W>
struct CpSmem {
  alignas(1024) uint8_t stages[W][2][kStageBytes];
  alignas(16) MergeSmem<W> merge;
};

// Warp-cooperative staging of one stage (nt tiles) into `dst` with the SWIZZLE_128B layout.
__device__ __forceinline__ void cp_stage(const DecodeParams& p, const WorkItem& w, int tile0, int nt,
                                         uint8_t* dst, int lane) {
  const int32_t* bt = p.block_table + static_cast<int64_t>(w.b) * p.bt_stride;
  const int c = lane & 15;       // 16-byte chunk of the 256-byte row
  const int rpar = lane >> 4;    // row parity handled by this half-warp
  for (int ti = 0; ti < nt; ++ti) {
    const int tile = tile0 + ti;
    const int64_t blk = __ldg(bt + tile / p.tiles_per_page);
    const int slot0 = (tile % p.tiles_per_page) * kTileTokens;
    const int64_t base_el = blk * p.cs_block + static_cast<int64_t>(w.kvh) * p.cs_head;
    uint8_t* kt = dst + ti * kKvTileBytesBf16 + (c >> 3) * kHalfBytes;
    uint8_t* vt = kt + kTileBytesBf16;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int r = 2 * it + rpar;
      const int64_t src_el = base_el + static_cast<int64_t>(slot0 + r) * p.cs_slot;
      const uint32_t off = swz_off(r, c & 7);
      cp_async_16(kt + off, p.k_cache + src_el * 2 + c * 16);
      cp_async_16(vt + off, p.v_cache + src_el * 2 + c * 16);
    }
  }
}

template <int W>
__global__ void __launch_bounds__(W * 32, 1) paged_decode_cpasync_kernel(const DecodeParams p) {
  extern __shared__ uint8_t smem_raw[];
  CpSmem<W>& sm =
      *reinterpret_cast<CpSmem<W>*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const WorkItem w = decode_item(p, blockIdx.x);
  if (!item_active(w)) return;  // uniform over the CTA
  const int ntiles = w.tile_end - w.tile_begin;
  const int nst = (ntiles + kStageTiles - 1) / kStageTiles;
  const LaneOffsets lo = make_lane_offsets(lane);
  uint32_t qf[8][2];
  load_q_frags(p, w, lane, qf);
  WarpState st;
  state_reset(st);

  if (warp < nst) {
    const int t0 = w.tile_begin + warp * kStageTiles;
    cp_stage(p, w, t0, min(kStageTiles, w.tile_end - t0), sm.stages[warp][0], lane);
  }
  cp_async_commit();
  int k = 0;
  for (int i = warp; i < nst; i += W, ++k) {
    const int tile0 = w.tile_begin + i * kStageTiles;
    const int nt = min(kStageTiles, w.tile_end - tile0);
    if (i + W < nst) {
      const int tn = tile0 + W * kStageTiles;
      cp_stage(p, w, tn, min(kStageTiles, w.tile_end - tn), sm.stages[warp][(k + 1) & 1], lane);
    }
    cp_async_commit();
    cp_async_wait<1>();
    __syncwarp();
    const int valid = min(nt * kTileTokens, w.seq_len - tile0 * kTileTokens);
    const bool need_mask = valid < nt * kTileTokens;
    uint8_t* stage = sm.stages[warp][k & 1];
    if (nt == 2)
      consume_stage_bf16<2>(stage, valid, need_mask, qf, lo, p.scale_log2, st, lane);
    else
      consume_stage_bf16<1>(stage, valid, need_mask, qf, lo, p.scale_log2, st, lane);
    __syncwarp();
  }
  cp_async_wait<0>();
  merge_and_store<W>(p, w, st, sm.merge, warp, lane, 1);
}

size_t decode_cpasync_smem_bytes() { return sizeof(CpSmem<kCpWarps>) + 1024; }

cudaError_t launch_decode_cpasync(const DecodeParams& p, cudaStream_t stream) {
  auto kern = paged_decode_cpasync_kernel<kCpWarps>;
  static thread_local int configured_dev = -1;
  const size_t smem = decode_cpasync_smem_bytes();
  if (cudaError_t e = ensure_dynamic_smem(kern, smem, &configured_dev)) return e;
  kern<<<p.total_items, kCpWarps * 32, smem, stream>>>(p);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// LSE reduce: out[b,h,:] = sum_s w_s o_s / sum_s w_s, w_s = exp(lse_s - max lse)
// One warp per (b, h) row, 4 rows per CTA; rows whose sequence used a single split were already
// written by the partial kernel and are skipped.
// ------------------------------------------------------------------------------------------------
constexpr int kReduceRowsPerCta = 4;

__global__ void __launch_bounds__(kReduceRowsPerCta * 32)
lse_reduce_kernel(__nv_bfloat16* __restrict__ out, int64_t out_stride_tok, float* __restrict__ lse,
                  const int32_t* __restrict__ seq_lens, int batch, int num_q_heads, int num_splits,
                  int tiles_per_split, const float* __restrict__ part_o,
                  const float* __restrict__ part_lse, int pdl) {
  if (pdl) griddep_wait();
  const int lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * kReduceRowsPerCta + (threadIdx.x >> 5);
  if (row >= static_cast<int64_t>(batch) * num_q_heads) return;
  const int b = static_cast<int>(row / num_q_heads);
  const int h = static_cast<int>(row % num_q_heads);
  const int seq_len = max(0, __ldg(seq_lens + b));
  const int tiles_total = (seq_len + kTileTokens - 1) / kTileTokens;
  const int splits = min(num_splits, (tiles_total + tiles_per_split - 1) / tiles_per_split);
  if (splits <= 1) return;

  const float* pl = part_lse + row * num_splits;
  const float* po = part_o + row * num_splits * kHeadDim;
  float mx = -CUDART_INF_F;
  for (int s = lane; s < splits; s += 32) mx = fmaxf(mx, pl[s]);
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  float denom = 0.f;
  if (mx != -CUDART_INF_F) {
    for (int s = 0; s < splits; ++s) {
      const float ls = pl[s];
      const float wgt = (ls == -CUDART_INF_F) ? 0.f : fast_exp2((ls - mx) * kLog2e);
      const float4 o = *reinterpret_cast<const float4*>(po + static_cast<int64_t>(s) * kHeadDim + lane * 4);
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
  *reinterpret_cast<uint2*>(out + static_cast<int64_t>(b) * out_stride_tok +
                            static_cast<int64_t>(h) * kHeadDim + lane * 4) = pk;
  if (lse != nullptr && lane == 0)
    lse[row] = denom > 0.f ? mx + logf(denom) : -CUDART_INF_F;
}

cudaError_t launch_lse_reduce(__nv_bfloat16* out, int64_t out_stride_tok, float* lse,
                              const int32_t* seq_lens, int batch, int num_q_heads, int num_splits,
                              int tiles_per_split, const float* part_o, const float* part_lse,
                              bool pdl, cudaStream_t stream) {
  const int64_t rows = static_cast<int64_t>(batch) * num_q_heads;
  const unsigned grid = static_cast<unsigned>((rows + kReduceRowsPerCta - 1) / kReduceRowsPerCta);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kReduceRowsPerCta * 32);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  const int pdl_i = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, lse_reduce_kernel, out, out_stride_tok, lse, seq_lens, batch,
                            num_q_heads, num_splits, tiles_per_split, part_o, part_lse, pdl_i);
}

}  // namespace b200
