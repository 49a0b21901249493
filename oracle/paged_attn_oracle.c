/*
 * paged_attn_oracle.c — plain-C CPU restatement of the paged-attention decode path.
 * TEST INFRASTRUCTURE, NOT PRODUCT CODE: only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.
 *
 * The reference (/root/reference, vllm-omni) holds no native code for this path; its arithmetic
 * lives in vLLM / FlashInfer / FlashAttention binaries (vLLM pinned 0.19.0, 0.22.0 installed).
 * Each function cites the vLLM file whose published semantics it follows. A second,
 * independent restatement (PyTorch fp32 SDPA) is oracle/oracle.py; the tests require both to agree.
 *
 * Build: make -C oracle   (gcc -O2 -fopenmp -shared)
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

/* ---- bf16 helpers ---- */
static inline float bf16_to_f32(uint16_t v) {
  uint32_t u = ((uint32_t)v) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

/* ---- e4m3 codec: OCP FP8 E4M3 (bias 7, max 448, NaN = S.1111.111, no inf) ---- */
float oracle_e4m3_to_f32(uint8_t c) {
  int s = c >> 7, e = (c >> 3) & 0xF, m = c & 7;
  float v;
  if (e == 0xF && m == 7) return NAN;
  if (e == 0) v = ldexpf((float)m, -9);
  else v = ldexpf(1.0f + (float)m / 8.0f, e - 7);
  return s ? -v : v;
}

/* round-to-nearest-even, saturate-to-finite: __nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3),
 * the conversion vLLM's reshape_and_cache_flash applies after `x / scale`
 * (semantics: vllm/v1/attention/ops/triton_reshape_and_cache_flash.py:96,110). */
uint8_t oracle_f32_to_e4m3_satfinite(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  uint8_t sign = (uint8_t)((u >> 31) << 7);
  uint32_t a = u & 0x7fffffffu;
  if (a > 0x7f800000u) return 0x7F;               /* NaN -> canonical NaN */
  if (a >= 0x43e80000u) return sign | 0x7E;       /* |x| >= 464 (incl. inf) -> 448; (448,464) rounds down below */
  int exp = (int)(a >> 23) - 127;
  if (exp < -6) {                                  /* subnormal target: multiples of 2^-9 */
    float af;
    memcpy(&af, &a, 4);
    float q = nearbyintf(af * 512.0f);             /* default rounding mode = RNE */
    return sign | (uint8_t)q;                      /* q == 8 is the first normal, code 0x08 */
  }
  uint32_t mant = a & 0x7fffffu;
  uint32_t m3 = mant >> 20, rem = mant & 0xfffffu;
  if (rem > 0x80000u || (rem == 0x80000u && (m3 & 1u))) m3++;
  if (m3 == 8) { m3 = 0; exp++; }
  if (exp > 8 || (exp == 8 && m3 > 6)) return sign | 0x7E;
  return sign | (uint8_t)(((exp + 7) << 3) | (int)m3);
}

void oracle_e4m3_encode(const float* x, uint8_t* out, int64_t n) {
  for (int64_t i = 0; i < n; ++i) out[i] = oracle_f32_to_e4m3_satfinite(x[i]);
}
void oracle_e4m3_decode(const uint8_t* c, float* out, int64_t n) {
  for (int64_t i = 0; i < n; ++i) out[i] = oracle_e4m3_to_f32(c[i]);
}

/*
 * KV-cache write (vllm/v1/attention/ops/triton_reshape_and_cache_flash.py:44-124).
 * key/value: bf16 [T,Hkv,d] with token/head strides (elements); caches: [nb,bs,Hkv,d] with
 * (block, slot, head) strides in cache elements; fp8 != 0 => u8 e4m3 cache.
 */
void oracle_reshape_and_cache(const uint16_t* key, const uint16_t* value, int64_t ks_tok,
                              int64_t ks_head, int64_t vs_tok, int64_t vs_head, void* k_cache,
                              void* v_cache, int64_t cs_block, int64_t cs_slot, int64_t cs_head,
                              const int64_t* slot_mapping, int32_t num_tokens, int32_t hkv,
                              int32_t d, int32_t bs, int32_t fp8, const float* k_scale,
                              const float* v_scale, int32_t scale_numel) {
  for (int32_t t = 0; t < num_tokens; ++t) {
    int64_t slot = slot_mapping[t];
    if (slot < 0) continue;                        /* :46-48 */
    int64_t blk = slot / bs, off = slot % bs;      /* :50-51 */
    for (int32_t h = 0; h < hkv; ++h) {
      const uint16_t* ksrc = key + t * ks_tok + h * ks_head;
      const uint16_t* vsrc = value + t * vs_tok + h * vs_head;
      int64_t dst = blk * cs_block + off * cs_slot + h * cs_head;
      if (!fp8) {
        memcpy((uint16_t*)k_cache + dst, ksrc, (size_t)d * 2);
        memcpy((uint16_t*)v_cache + dst, vsrc, (size_t)d * 2);
      } else {
        float ksc = k_scale[scale_numel > 1 ? h : 0], vsc = v_scale[scale_numel > 1 ? h : 0];
        for (int32_t i = 0; i < d; ++i) {
          ((uint8_t*)k_cache)[dst + i] = oracle_f32_to_e4m3_satfinite(bf16_to_f32(ksrc[i]) / ksc);
          ((uint8_t*)v_cache)[dst + i] = oracle_f32_to_e4m3_satfinite(bf16_to_f32(vsrc[i]) / vsc);
        }
      }
    }
  }
}

/*
 * Paged decode attention, fp32 in / double accumulate. One (sequence, query head) per OpenMP task.
 * Semantics of flash_attn_varlen_func with block_table + seqused_k at max_seqlen_q = 1
 * (vllm/v1/attention/backends/flash_attn.py:758-819); GQA as in cpu_attn.py:449 (enable_gqa).
 * q: bf16 [B,Hq,d] (token stride given); out: f32 [B,Hq,d]; lse: f32 [B,Hq] natural log.
 * scratch: f32 [num_threads, max_seq_len] is avoided by a two-pass (max, then sum) evaluation.
 */
void oracle_paged_decode(const uint16_t* q, int64_t q_stride_tok, const void* k_cache,
                         const void* v_cache, int64_t cs_block, int64_t cs_slot, int64_t cs_head,
                         float* out, float* lse, const int32_t* block_table, int64_t bt_stride,
                         const int32_t* seq_lens, int32_t batch, int32_t hq, int32_t hkv, int32_t d,
                         int32_t bs, float softmax_scale, int32_t fp8, const float* k_scale,
                         const float* v_scale, int32_t scale_numel) {
  const int32_t group = hq / hkv;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
  for (int32_t b = 0; b < batch; ++b) {
    for (int32_t h = 0; h < hq; ++h) {
      const int32_t S = seq_lens[b];
      float* o = out + ((int64_t)b * hq + h) * d;
      for (int32_t i = 0; i < d; ++i) o[i] = 0.f;
      if (S <= 0) { if (lse) lse[(int64_t)b * hq + h] = -INFINITY; continue; }
      const int32_t kvh = h / group;
      const float ksc = fp8 ? k_scale[scale_numel > 1 ? kvh : 0] : 1.f;
      const float vsc = fp8 ? v_scale[scale_numel > 1 ? kvh : 0] : 1.f;
      const uint16_t* qr = q + (int64_t)b * q_stride_tok + (int64_t)h * d;
      float qf[512];
      for (int32_t i = 0; i < d; ++i) qf[i] = bf16_to_f32(qr[i]);
      /* pass 1: row max */
      double mx = -INFINITY;
      for (int32_t j = 0; j < S; ++j) {
        int64_t blk = block_table[(int64_t)b * bt_stride + j / bs];
        int64_t base = blk * cs_block + (int64_t)(j % bs) * cs_slot + (int64_t)kvh * cs_head;
        double s = 0.0;
        if (fp8) { const uint8_t* kr = (const uint8_t*)k_cache + base;
          for (int32_t i = 0; i < d; ++i) s += (double)qf[i] * (double)(oracle_e4m3_to_f32(kr[i]) * ksc);
        } else { const uint16_t* kr = (const uint16_t*)k_cache + base;
          for (int32_t i = 0; i < d; ++i) s += (double)qf[i] * (double)bf16_to_f32(kr[i]); }
        s *= softmax_scale;
        if (s > mx) mx = s;
      }
      /* pass 2: exp-sum and weighted V */
      double den = 0.0;
      double acc[512];
      for (int32_t i = 0; i < d; ++i) acc[i] = 0.0;
      for (int32_t j = 0; j < S; ++j) {
        int64_t blk = block_table[(int64_t)b * bt_stride + j / bs];
        int64_t base = blk * cs_block + (int64_t)(j % bs) * cs_slot + (int64_t)kvh * cs_head;
        double s = 0.0;
        if (fp8) { const uint8_t* kr = (const uint8_t*)k_cache + base;
          for (int32_t i = 0; i < d; ++i) s += (double)qf[i] * (double)(oracle_e4m3_to_f32(kr[i]) * ksc);
        } else { const uint16_t* kr = (const uint16_t*)k_cache + base;
          for (int32_t i = 0; i < d; ++i) s += (double)qf[i] * (double)bf16_to_f32(kr[i]); }
        double p = exp(s * softmax_scale - mx);
        den += p;
        if (fp8) { const uint8_t* vr = (const uint8_t*)v_cache + base;
          for (int32_t i = 0; i < d; ++i) acc[i] += p * (double)(oracle_e4m3_to_f32(vr[i]) * vsc);
        } else { const uint16_t* vr = (const uint16_t*)v_cache + base;
          for (int32_t i = 0; i < d; ++i) acc[i] += p * (double)bf16_to_f32(vr[i]); }
      }
      for (int32_t i = 0; i < d; ++i) o[i] = (float)(acc[i] / den);
      if (lse) lse[(int64_t)b * hq + h] = (float)(mx + log(den));
    }
  }
}
