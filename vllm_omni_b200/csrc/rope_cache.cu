// Fused rotary embedding (in place on q and k) + KV-cache write for sm_100a.
//
// Replaces the pair of launches the reference's attention layers issue before every attention call
// (vllm_omni/model_executor/models/qwen2_5_omni/qwen2_old.py:151-154: `q, k = self.rotary_emb(positions,
// q, k)` then `self.attn(q, k, v)` -> unified_kv_cache_update) with one, behind vLLM's own hook
// `AttentionImpl.do_rope_and_kv_cache_update` (vllm/v1/attention/backend.py:815-840): K makes one
// round trip through HBM instead of two and a launch disappears from every layer.
//
// Arithmetic = vLLM's RotaryEmbedding / MRotaryEmbedding in the activation dtype (bf16), restated
// from vllm/model_executor/layers/rotary_embedding/common.py:144-174 and mrope.py:263-322:
//     o1 = bf16(bf16(x1 * cos) - bf16(x2 * sin)),   o2 = bf16(bf16(x2 * cos) + bf16(x1 * sin))
// with (x1, x2) = (x[p], x[p + rot/2]) for the NeoX layout and (x[2p], x[2p + 1]) for GPT-J, cos / sin
// = cos_sin_cache[pos][p] / [pos][rot/2 + p]. Multimodal M-RoPE (positions [3, T]): pair p takes its
// position from row r(p) — chunked sections (Qwen2.5-Omni) or the interleaved T/H/W pattern
// (Qwen3-Omni, mrope.py:190-198). The rotated K row (rounded to bf16, as the unfused pair of kernels
// leaves it) is what the cache write stores / quantises; elements past rot_dim pass through.
#include <cuda_fp8.h>

#include "kernels.h"

namespace b200 {

__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// o1, o2 of one pair with every operation rounded to bf16 (bf16 tensor arithmetic).
__device__ __forceinline__ void rope_pair(float x1, float x2, float c, float s, float& o1, float& o2) {
  o1 = bf16_round(bf16_round(x1 * c) - bf16_round(x2 * s));
  o2 = bf16_round(bf16_round(x2 * c) + bf16_round(x1 * s));
}

__device__ __forceinline__ void unpack8(const int4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ int4 pack8(const float (&f)[8]) {
  int4 v;
  uint32_t* w = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) w[i] = pack_bf16x2(f[2 * i], f[2 * i + 1]);
  return v;
}

// Store 8 consecutive elements (already bf16-representable floats) of a K/V row into the cache.
template <bool FP8>
__device__ __forceinline__ void cache_store8(uint8_t* cache, int64_t dst_el, const float (&f)[8], float scale) {
  if constexpr (!FP8) {
    *reinterpret_cast<int4*>(cache + dst_el * 2) = pack8(f);
  } else {
    uint32_t w[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      uint32_t q[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) q[i] = __nv_cvt_float_to_fp8(__fdiv_rn(f[4 * j + i], scale), __NV_SATFINITE, __NV_E4M3);
      w[j] = q[0] | (q[1] << 8) | (q[2] << 16) | (q[3] << 24);
    }
    *reinterpret_cast<uint2*>(cache + dst_el) = make_uint2(w[0], w[1]);
  }
}

// Work item = 8 rotation pairs (16 elements) or 8 pass-through elements of one (token, head) row.
template <bool FP8>
__global__ void __launch_bounds__(256) rope_cache_kernel(const RopeCacheParams p) {
  griddep_wait();
  griddep_launch_dependents();
  const int half = p.rot_dim / 2;
  const int rot_items = half / 8;                       // items that rotate 8 pairs
  const int pass_items = (p.head_size - p.rot_dim) / 8;  // items that copy 8 elements
  const int qk_items = rot_items + pass_items;
  const int v_items = p.head_size / 8;
  const int64_t per_tok = static_cast<int64_t>(p.num_q_heads + p.num_kv_heads) * qk_items +
                          static_cast<int64_t>(p.num_kv_heads) * v_items;
  const int64_t total = per_tok * p.num_tokens;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t t = i / per_tok;
    int r = static_cast<int>(i % per_tok);
    // ---- which row of the token, and which item of the row
    int kind, h, item;  // kind 0 = q, 1 = k, 2 = v
    if (r < p.num_q_heads * qk_items) {
      kind = 0, h = r / qk_items, item = r % qk_items;
    } else if ((r -= p.num_q_heads * qk_items) < p.num_kv_heads * qk_items) {
      kind = 1, h = r / qk_items, item = r % qk_items;
    } else {
      r -= p.num_kv_heads * qk_items;
      kind = 2, h = r / v_items, item = r % v_items;
    }
    // ---- cache destination of this token (k, v rows only)
    int64_t dst_row = -1;
    if (kind != 0 && t < p.num_slot_tokens) {
      const int64_t slot = __ldg(p.slot_mapping + t);
      const int64_t blk = slot / p.block_size;
      if (slot >= 0 && blk < p.num_blocks)
        dst_row = blk * p.cs_block + (slot % p.block_size) * p.cs_slot + static_cast<int64_t>(h) * p.cs_head;
    }
    if (kind == 2) {
      if (dst_row < 0) continue;
      const int4 v = ldg_nc_v4(p.value + t * p.value_stride_tok + static_cast<int64_t>(h) * p.value_stride_head + item * 8);
      float f[8];
      unpack8(v, f);
      cache_store8<FP8>(p.v_cache, dst_row + item * 8, f, FP8 ? __ldg(p.v_scale + h * p.scale_stride) : 1.f);
      continue;
    }
    __nv_bfloat16* row = kind == 0 ? p.query + t * p.q_stride_tok + static_cast<int64_t>(h) * p.head_size
                                   : p.key + t * p.key_stride_tok + static_cast<int64_t>(h) * p.key_stride_head;
    const float kscale = (FP8 && kind == 1) ? __ldg(p.k_scale + h * p.scale_stride) : 1.f;
    if (item >= rot_items) {  // pass-through tail of a partially rotated head
      if (kind == 1 && dst_row >= 0) {
        const int e0 = p.rot_dim + (item - rot_items) * 8;
        float f[8];
        unpack8(*reinterpret_cast<const int4*>(row + e0), f);
        cache_store8<FP8>(p.k_cache, dst_row + e0, f, kscale);
      }
      continue;
    }
    const int p0 = item * 8;  // first pair of the item
    // ---- cos / sin of the 8 pairs
    float c[8], s[8];
    if (p.pos_rows == 1) {
      const __nv_bfloat16* cs = p.cos_sin + __ldg(p.positions + t) * p.cos_sin_stride;
      unpack8(ldg_nc_v4(cs + p0), c);
      unpack8(ldg_nc_v4(cs + half + p0), s);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int pr = p0 + j;
        int rowsel;
        if (p.mrope_interleaved) {
          const int m = pr % 3;
          rowsel = (m == 1 && pr < p.mrope_section[1] * 3) ? 1 : ((m == 2 && pr < p.mrope_section[2] * 3) ? 2 : 0);
        } else {
          rowsel = pr < p.mrope_section[0] ? 0 : (pr < p.mrope_section[0] + p.mrope_section[1] ? 1 : 2);
        }
        const __nv_bfloat16* cs = p.cos_sin + __ldg(p.positions + rowsel * p.pos_stride_row + t) * p.cos_sin_stride;
        c[j] = __bfloat162float(cs[pr]);
        s[j] = __bfloat162float(cs[half + pr]);
      }
    }
    // ---- rotate
    float x1[8], x2[8], o1[8], o2[8];
    if (p.is_neox) {
      unpack8(*reinterpret_cast<const int4*>(row + p0), x1);
      unpack8(*reinterpret_cast<const int4*>(row + half + p0), x2);
    } else {
      float a[8], b[8];
      unpack8(*reinterpret_cast<const int4*>(row + 2 * p0), a);
      unpack8(*reinterpret_cast<const int4*>(row + 2 * p0 + 8), b);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        x1[j] = a[2 * j], x2[j] = a[2 * j + 1];
        x1[4 + j] = b[2 * j], x2[4 + j] = b[2 * j + 1];
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) rope_pair(x1[j], x2[j], c[j], s[j], o1[j], o2[j]);
    if (p.is_neox) {
      *reinterpret_cast<int4*>(row + p0) = pack8(o1);
      *reinterpret_cast<int4*>(row + half + p0) = pack8(o2);
      if (kind == 1 && dst_row >= 0) {
        cache_store8<FP8>(p.k_cache, dst_row + p0, o1, kscale);
        cache_store8<FP8>(p.k_cache, dst_row + half + p0, o2, kscale);
      }
    } else {
      float a[8], b[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        a[2 * j] = o1[j], a[2 * j + 1] = o2[j];
        b[2 * j] = o1[4 + j], b[2 * j + 1] = o2[4 + j];
      }
      *reinterpret_cast<int4*>(row + 2 * p0) = pack8(a);
      *reinterpret_cast<int4*>(row + 2 * p0 + 8) = pack8(b);
      if (kind == 1 && dst_row >= 0) {
        cache_store8<FP8>(p.k_cache, dst_row + 2 * p0, a, kscale);
        cache_store8<FP8>(p.k_cache, dst_row + 2 * p0 + 8, b, kscale);
      }
    }
  }
}

cudaError_t launch_rope_cache(const RopeCacheParams& p, bool fp8, int num_sms, bool pdl, cudaStream_t stream) {
  const int qk_items = p.rot_dim / 16 + (p.head_size - p.rot_dim) / 8;
  const int64_t per_tok = static_cast<int64_t>(p.num_q_heads + p.num_kv_heads) * qk_items +
                          static_cast<int64_t>(p.num_kv_heads) * (p.head_size / 8);
  const int64_t total = per_tok * p.num_tokens;
  if (total == 0) return cudaSuccess;
  const int threads = 256;
  int64_t blocks = (total + threads - 1) / threads;
  const int64_t cap = static_cast<int64_t>(num_sms) * 8;
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
  return fp8 ? cudaLaunchKernelEx(&cfg, rope_cache_kernel<true>, p) : cudaLaunchKernelEx(&cfg, rope_cache_kernel<false>, p);
}

}  // namespace b200
