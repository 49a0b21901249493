// KV-cache write (reshape_and_cache) and KV gather for sm_100a.
//
// Cache write: replaces vLLM `reshape_and_cache_flash`
// (vllm/v1/attention/ops/triton_reshape_and_cache_flash.py:44-124 restates its semantics:
//  negative slot skip :46-48, block/offset split :50-51, division by the scale :96,110).
// Work unit = one (token, kv head, K|V) row of head_size elements; 16 lanes move one
// 128-element row with 16-byte loads (bf16) and 16- or 8-byte stores (bf16 / e4m3).
#include <cuda_fp8.h>

#include "kernels.h"

namespace b200 {

template <bool FP8>
__global__ void __launch_bounds__(256) cache_write_kernel(const CacheWriteParams p) {
  // programmatic dependent launch: this grid may be resident before the previous kernel on the
  // stream has finished; nothing is read or written before the wait, and the kernel after this one
  // (the decode) may start its own prologue right away
  griddep_wait();
  griddep_launch_dependents();
  const int lanes_per_row = p.head_size / 8;  // 8 bf16 = 16 B per lane; 16 for head_size 128
  const int64_t rows = static_cast<int64_t>(p.num_tokens) * p.num_kv_heads * 2;
  const int64_t total = rows * lanes_per_row;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % lanes_per_row);
    int64_t r = i / lanes_per_row;
    const int is_v = static_cast<int>(r & 1);
    r >>= 1;
    const int h = static_cast<int>(r % p.num_kv_heads);
    const int64_t t = r / p.num_kv_heads;
    const int64_t slot = __ldg(p.slot_mapping + t);
    if (slot < 0) continue;  // padding token
    const int64_t blk = slot / p.block_size;
    const int off = static_cast<int>(slot % p.block_size);
    if (blk >= p.num_blocks) continue;  // never write outside the cache
    const __nv_bfloat16* src = is_v ? p.value + t * p.value_stride_tok + h * p.value_stride_head
                                    : p.key + t * p.key_stride_tok + h * p.key_stride_head;
    const int4 v = ldg_nc_v4(src + c * 8);
    uint8_t* cache = is_v ? p.v_cache : p.k_cache;
    const int64_t dst_el = blk * p.cs_block + static_cast<int64_t>(off) * p.cs_slot +
                           static_cast<int64_t>(h) * p.cs_head + c * 8;
    if constexpr (!FP8) {
      *reinterpret_cast<int4*>(cache + dst_el * 2) = v;
    } else {
      const float scale = __ldg((is_v ? p.v_scale : p.k_scale) + h * p.scale_stride);
      const __nv_bfloat162* xs = reinterpret_cast<const __nv_bfloat162*>(&v);
      uint32_t w[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float2 a = __bfloat1622float2(xs[2 * j]);
        const float2 b = __bfloat1622float2(xs[2 * j + 1]);
        // IEEE division then RNE + saturate-to-finite e4m3: the same two steps as vLLM's
        // fp8::scaled_convert (x / scale -> __nv_cvt_float_to_fp8(…, __NV_SATFINITE, __NV_E4M3)).
        const uint32_t q0 = __nv_cvt_float_to_fp8(__fdiv_rn(a.x, scale), __NV_SATFINITE, __NV_E4M3);
        const uint32_t q1 = __nv_cvt_float_to_fp8(__fdiv_rn(a.y, scale), __NV_SATFINITE, __NV_E4M3);
        const uint32_t q2 = __nv_cvt_float_to_fp8(__fdiv_rn(b.x, scale), __NV_SATFINITE, __NV_E4M3);
        const uint32_t q3 = __nv_cvt_float_to_fp8(__fdiv_rn(b.y, scale), __NV_SATFINITE, __NV_E4M3);
        w[j] = q0 | (q1 << 8) | (q2 << 16) | (q3 << 24);
      }
      *reinterpret_cast<uint2*>(cache + dst_el) = make_uint2(w[0], w[1]);
    }
  }
}

cudaError_t launch_cache_write(const CacheWriteParams& p, bool fp8, int num_sms, bool pdl,
                               cudaStream_t stream) {
  const int64_t total =
      static_cast<int64_t>(p.num_tokens) * p.num_kv_heads * 2 * (p.head_size / 8);
  if (total == 0) return cudaSuccess;
  const int threads = 256;
  int64_t blocks = (total + threads - 1) / threads;
  const int64_t cap = static_cast<int64_t>(num_sms) * 8;  // whole waves of resident CTAs
  if (blocks > cap) blocks = cap;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(blocks));
  cfg.blockDim = dim3(threads);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return fp8 ? cudaLaunchKernelEx(&cfg, cache_write_kernel<true>, p)
             : cudaLaunchKernelEx(&cfg, cache_write_kernel<false>, p);
}

// Gather of one request's rows: dst[s, h, :] = cache[block_ids[s / bs], s % bs, h, :], s < seq_len.
// Device twin of kv_transfer_manager.py:829-846 in the reference.
__global__ void __launch_bounds__(256) gather_kv_kernel(const GatherParams p) {
  const int row_bytes = p.head_size * p.elem_bytes;
  const int chunks = row_bytes / 16;
  const int64_t total = static_cast<int64_t>(p.seq_len) * p.num_kv_heads * chunks;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % chunks);
    int64_t r = i / chunks;
    const int h = static_cast<int>(r % p.num_kv_heads);
    const int64_t s = r / p.num_kv_heads;
    const int page = static_cast<int>(s / p.block_size);
    if (page >= p.num_block_ids) continue;
    const int64_t blk = __ldg(p.block_ids + page);
    if (blk < 0 || blk >= p.num_blocks) continue;
    const int64_t src_el = blk * p.cs_block + (s % p.block_size) * p.cs_slot +
                           static_cast<int64_t>(h) * p.cs_head;
    const int4 v = ldg_nc_v4(p.cache + src_el * p.elem_bytes + c * 16);
    *reinterpret_cast<int4*>(p.dst + (s * p.num_kv_heads + h) * row_bytes + c * 16) = v;
  }
}

cudaError_t launch_gather_kv(const GatherParams& p, int num_sms, cudaStream_t stream) {
  const int64_t total = static_cast<int64_t>(p.seq_len) * p.num_kv_heads *
                        (p.head_size * p.elem_bytes / 16);
  if (total == 0) return cudaSuccess;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = static_cast<int64_t>(num_sms) * 8;
  if (blocks > cap) blocks = cap;
  gather_kv_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(p);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// FP8 prefill staging: the e4m3 pages the prefill sequences of a step own are rewritten ONCE per layer as bf16 pages
// (the e4m3 values themselves, exact in bf16; the dequantisation scales stay with the attention kernel) so that the
// tcgen05 kernel of paged_prefill_umma.cu serves fp8 caches too. 5 bytes of HBM traffic per cached element against
// 4 * d * Hq / Hkv flops per (query, key) pair: 1-2 % of a prefill step, and the staged pages are read back from L2.
//
// Staged layout: page (b, j) of sequence b (j < ceil(S_b / bs)) is staged page cum[b] + j, cum = exclusive prefix
// of pages per sequence; a staged page is [Hkv][K|V][bs][d] bf16 like the cache itself (layout.py). Tokens at or
// beyond S_b are staged as zeros (what the unused slots of a cache page hold is undefined).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) stage_fp8_plan_kernel(const StageFp8Params p) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < p.batch; base += 1024) {
    const int b = base + tid;
    const int n = b < p.batch ? (max(0, __ldg(p.seq_lens + b)) + p.block_size - 1) / p.block_size : 0;
    int x = n;  // inclusive scan inside the warp
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, off);
      if (lane >= off) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = s_warp[lane];
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, w, off);
        if (lane >= off) w += y;
      }
      s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const int carry = s_carry;
    const int excl = carry + (warp > 0 ? s_warp[warp - 1] : 0) + x - n;
    if (b < p.batch) p.cum[b] = excl;
    __syncthreads();
    if (tid == 1023) s_carry = carry + s_warp[31];
    __syncthreads();
  }
  if (tid == 0) p.cum[p.batch] = s_carry;
}

constexpr int kStagePagesPerCta = 4;

__global__ void __launch_bounds__(256) stage_fp8_pages_kernel(const StageFp8Params p) {
  const int b = blockIdx.y;
  const int S = max(0, __ldg(p.seq_lens + b));
  const int npages = (S + p.block_size - 1) / p.block_size;
  const int j0 = blockIdx.x * kStagePagesPerCta;
  if (j0 >= npages) return;
  const int64_t base = p.cum[b];
  const int units_per_row = kHeadDim / 16;                              // 16 e4m3 bytes -> 32 bf16 bytes
  const int units_per_page = p.num_kv_heads * 2 * p.block_size * units_per_row;
  const int64_t page_elems = static_cast<int64_t>(p.num_kv_heads) * 2 * p.block_size * kHeadDim;
  const int j1 = min(npages, j0 + kStagePagesPerCta);
  for (int j = j0; j < j1; ++j) {
    const int64_t dst_page = base + j;
    if (dst_page >= p.capacity_pages) return;  // the host sized the staging for the step: never taken
    const int64_t blk = __ldg(p.block_table + b * p.bt_stride + j);
    if (threadIdx.x == 0) p.stage_bt[static_cast<int64_t>(b) * p.max_pages + j] = static_cast<int32_t>(dst_page);
    const bool blk_ok = blk >= 0 && blk < p.num_blocks;
    __nv_bfloat16* dst = p.pages + dst_page * page_elems;
    for (int u = threadIdx.x; u < units_per_page; u += blockDim.x) {
      const int c = u % units_per_row;
      int r = u / units_per_row;
      const int slot = r % p.block_size;
      r /= p.block_size;
      const int is_v = r & 1;
      const int h = r >> 1;
      int4 raw = make_int4(0, 0, 0, 0);
      if (blk_ok && j * p.block_size + slot < S) {
        const uint8_t* src = (is_v ? p.v_cache : p.k_cache) + blk * p.cs_block + static_cast<int64_t>(slot) * p.cs_slot +
                             static_cast<int64_t>(h) * p.cs_head + c * 16;
        raw = ldg_nc_v4(src);
      }
      const uint32_t w[4] = {static_cast<uint32_t>(raw.x), static_cast<uint32_t>(raw.y), static_cast<uint32_t>(raw.z),
                             static_cast<uint32_t>(raw.w)};
      uint32_t o[8];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const __half2_raw hr = __nv_cvt_fp8x2_to_halfraw2(
              static_cast<__nv_fp8x2_storage_t>((w[k] >> (16 * half)) & 0xffffu), __NV_E4M3);
          const float2 f = __half22float2(__half2(hr));
          o[2 * k + half] = pack_bf16x2(f.x, f.y);  // exact: an e4m3 value has 4 significant bits
        }
      }
      uint4* d = reinterpret_cast<uint4*>(dst + ((static_cast<int64_t>(h) * 2 + is_v) * p.block_size + slot) * kHeadDim +
                                          c * 16);
      d[0] = make_uint4(o[0], o[1], o[2], o[3]);
      d[1] = make_uint4(o[4], o[5], o[6], o[7]);
    }
  }
}

cudaError_t launch_stage_fp8(const StageFp8Params& p, cudaStream_t stream) {
  if (p.batch <= 0 || p.max_pages <= 0) return cudaSuccess;
  stage_fp8_plan_kernel<<<1, 1024, 0, stream>>>(p);
  if (cudaError_t e = cudaGetLastError()) return e;
  const dim3 grid(static_cast<unsigned>((p.max_pages + kStagePagesPerCta - 1) / kStagePagesPerCta),
                  static_cast<unsigned>(p.batch));
  stage_fp8_pages_kernel<<<grid, 256, 0, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace b200
