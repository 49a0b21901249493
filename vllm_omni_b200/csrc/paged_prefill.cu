// Causal varlen attention over the paged KV cache for query length >= 1 (prefill, chunked
// prefill / extend, mixed batches) on sm_100a — the path a registered backend needs for the
// non-decode tokens of a step (vllm/v1/attention/backends/flash_attn.py:758-819 with
// max_query_len > 1). The new tokens' K/V are already in the cache (do_kv_cache_update runs first).
//
// Work unit = (sequence, 16-query-token tile, kv head [, 8-head sub-group]): one CTA whose warp w
// owns query head (kv head * G + w) for the tile's 16 tokens, so the G heads of a GQA group share
// every KV tile that is staged — KV traffic per CTA is 1/G of a head-per-CTA layout. A producer
// warp streams the sequence's pages through a ring of stages (two 16-token K tiles + V tiles per
// stage) with the same 5-D / 4-D SWIZZLE_128B tensor maps as the decode kernel; consumer warps
// run flash-attention style online softmax on mma.sync.m16n8k16 (S = Q K^T with Q as the
// register-resident A operand, O += P V with P re-used from the C fragments), fp32 accumulate,
// log2-domain exponentials, lazily refreshed reference max. FP8-e4m3 caches are converted to f16
// in registers (exact) and multiplied with f16 MMAs; k_scale folds into the softmax scale and
// v_scale into the final normalisation.
//
// Causality: query i of a sequence with q_len new tokens and S cached tokens sits at position
// S - q_len + i and attends to cache positions <= its own. A tile therefore streams KV tokens
// [0, min(S, S - q_len + i0 + 16)); only its last one or two KV tiles need the mask.
#include "kernels.h"
#include "sk_math.cuh"

namespace b200 {

constexpr int kPfTok = 16;     // query tokens per CTA
constexpr int kPfKt = 2;       // 16-token KV tiles per stage
constexpr int kPfStages = 4;   // ring depth
constexpr int kPfWarps = 8;    // consumer warps (= query heads per CTA); + 1 producer warp

template <int SB>
struct PfSmem {
  alignas(1024) uint8_t ring[kPfStages][kPfKt][SB];  // SB = K tile + V tile bytes
  alignas(8) uint64_t full[kPfStages];
  alignas(8) uint64_t empty[kPfStages];
};

struct PfLane {
  uint32_t k[4];  // K as B operand: ldmatrix.x4 address for k-step mod 4 (within a 2 KiB half)
  uint32_t v[4];  // V as B operand: ldmatrix.x4.trans address for d-tile pair mod 4 (within a half)
};
__device__ __forceinline__ PfLane pf_make_lane(int lane) {
  PfLane o;
  const int m = lane >> 3, r = lane & 7;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    // K: matrices (tok 0-7, c) (tok 0-7, c+1) (tok 8-15, c) (tok 8-15, c+1), c = 2i
    o.k[i] = swz_off(r + 8 * (m >> 1), 2 * i + (m & 1));
    // V: matrices (tok 0-7, c) (tok 8-15, c) (tok 0-7, c+1) (tok 8-15, c+1), c = 2i, transposed
    o.v[i] = swz_off(r + 8 * (m & 1), 2 * i + (m >> 1));
  }
  return o;
}

// CTA -> (sequence, query tile). Sequence b owns tile slots [qsl[b] / 16 + b, qsl[b+1] / 16 + b + 1),
// which always holds its ceil(q_len / 16) tiles; unused slots exit.
__device__ __forceinline__ bool pf_locate(const PrefillParams& p, int slot, int& b, int& qt, int& q0, int& q_len) {
  int lo = 0, hi = p.batch;  // start(lo) <= slot < start(hi)
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if ((__ldg(p.query_start_loc + mid) >> 4) + mid <= slot) lo = mid; else hi = mid;
  }
  b = lo;
  q0 = __ldg(p.query_start_loc + b);
  q_len = __ldg(p.query_start_loc + b + 1) - q0;
  qt = slot - ((q0 >> 4) + b);
  return qt * kPfTok < q_len;
}

template <bool FP8>
__global__ void __launch_bounds__((kPfWarps + 1) * 32, 1)
paged_prefill_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                     const PrefillParams p) {
  constexpr int TB = FP8 ? kTileBytesFp8 : kTileBytesBf16;  // one K (or V) tile
  constexpr int SB = 2 * TB;
  extern __shared__ uint8_t smem_raw[];
  PfSmem<SB>& sm = *reinterpret_cast<PfSmem<SB>*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  int b, qt, q0, q_len;
  // heavy (late) tiles first: they stream the most KV
  if (!pf_locate(p, static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x), b, qt, q0, q_len)) return;
  const int hs = blockIdx.y;
  const int kvh = hs / p.subgroups, sub = hs % p.subgroups;
  const int heads_here = min(kPfWarps, p.group - sub * kPfWarps);
  const int S = max(0, __ldg(p.seq_lens + b));
  const int i0 = qt * kPfTok;
  const int ctx0 = S - q_len;  // cache position of the sequence's first new token
  const int kv_end = max(0, min(S, ctx0 + i0 + kPfTok));
  const int ntiles = (kv_end + kTileTokens - 1) / kTileTokens;
  const int nstages = (ntiles + kPfKt - 1) / kPfKt;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kPfStages; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], heads_here);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
  }
  __syncthreads();

  if (warp == kPfWarps) {
    // ===================== producer warp =====================
    const uint64_t pol = l2_policy_evict_last();  // the other query tiles of the sequence re-read these pages
    int page0 = -32, ids = 0;
    for (int st = 0; st < nstages; ++st) {
      const int slot = st % kPfStages;
      if (st >= kPfStages) mbar_wait(&sm.empty[slot], ((st / kPfStages) - 1) & 1);
      const int t0 = st * kPfKt;
      const int nt = min(kPfKt, ntiles - t0);
      if (lane == 0) mbar_arrive_expect_tx(&sm.full[slot], static_cast<uint32_t>(nt * SB));
#pragma unroll
      for (int ti = 0; ti < kPfKt; ++ti) {
        if (ti >= nt) break;
        const int tile = t0 + ti;
        const int page = tile >> p.tpp_shift;
        if (page >= page0 + 32) {
          page0 = page & ~31;
          const int pg = page0 + lane;
          ids = (pg << p.tpp_shift) < ntiles ? __ldg(p.block_table + static_cast<int64_t>(b) * p.bt_stride + pg) : 0;
        }
        const int blk = __shfl_sync(0xffffffffu, ids, page - page0);
        if (lane == 0) {
          const uint32_t dst = smem_u32(sm.ring[slot][ti]), bar = smem_u32(&sm.full[slot]);
          const int slot0 = (tile & p.tpp_mask) * kTileTokens;
          if (FP8) {
            tma_load_4d_s(dst, &tm_k, bar, 0, slot0, kvh, blk, pol);
            tma_load_4d_s(dst + TB, &tm_v, bar, 0, slot0, kvh, blk, pol);
          } else {
            tma_load_5d(dst, &tm_k, bar, 0, slot0, 0, kvh, blk, pol);
            tma_load_5d(dst + TB, &tm_v, bar, 0, slot0, 0, kvh, blk, pol);
          }
        }
      }
    }
    return;
  }
  if (warp >= heads_here) return;

  // ===================== consumer warps: one query head x 16 tokens each =====================
  const int g = lane >> 2, t = lane & 3;
  const int h = kvh * p.group + sub * kPfWarps + warp;
  const PfLane pl = pf_make_lane(lane);
  const bool row0_ok = i0 + g < q_len, row1_ok = i0 + g + 8 < q_len;
  // Q as A operand: qa[kk] = {Q[g][lo], Q[g+8][lo], Q[g][hi], Q[g+8][hi]} of k-step kk.
  // bf16: lo = d 16kk+2t,+1, hi = d 16kk+2t+8,+9.  fp8: f16, lo = d 16kk+4t,+1, hi = d 16kk+4t+2,+3
  // (the k permutation the e4m3 K fragments come in, see sk_math.cuh).
  uint32_t qa[8][4];
  {
    const uint32_t* r0 = reinterpret_cast<const uint32_t*>(p.q + static_cast<int64_t>(q0 + i0 + g) * p.q_stride_tok +
                                                           static_cast<int64_t>(h) * kHeadDim);
    const uint32_t* r1 = reinterpret_cast<const uint32_t*>(p.q + static_cast<int64_t>(q0 + i0 + g + 8) * p.q_stride_tok +
                                                           static_cast<int64_t>(h) * kHeadDim);
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) {
      if (FP8) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const uint32_t x0 = row0_ok ? __ldg(r0 + kk * 8 + 2 * t + i) : 0u;
          const uint32_t x1 = row1_ok ? __ldg(r1 + kk * 8 + 2 * t + i) : 0u;
          qa[kk][2 * i] = pack_f16x2(__uint_as_float(x0 << 16), __uint_as_float(x0 & 0xffff0000u));
          qa[kk][2 * i + 1] = pack_f16x2(__uint_as_float(x1 << 16), __uint_as_float(x1 & 0xffff0000u));
        }
      } else {
        qa[kk][0] = row0_ok ? __ldg(r0 + kk * 8 + t) : 0u;
        qa[kk][1] = row1_ok ? __ldg(r1 + kk * 8 + t) : 0u;
        qa[kk][2] = row0_ok ? __ldg(r0 + kk * 8 + t + 4) : 0u;
        qa[kk][3] = row1_ok ? __ldg(r1 + kk * 8 + t + 4) : 0u;
      }
    }
  }
  float scale = p.scale_log2;
  if (FP8) scale *= __ldg(p.k_scale + kvh * p.scale_stride);
  const int pos0 = ctx0 + i0 + g, pos1 = pos0 + 8;  // cache positions of this thread's two rows

  float m0 = -CUDART_INF_F, m1 = -CUDART_INF_F, l0 = 0.f, l1 = 0.f;
  float o[16][4];
#pragma unroll
  for (int j = 0; j < 16; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;

  for (int st = 0; st < nstages; ++st) {
    const int slot = st % kPfStages;
    mbar_wait(&sm.full[slot], (st / kPfStages) & 1);
    const int nt = min(kPfKt, ntiles - st * kPfKt);
    // ---- S = Q K^T for the stage's tiles: s[ti][n-tile][c]
    float s[kPfKt][2][4];
#pragma unroll
    for (int ti = 0; ti < kPfKt; ++ti) {
#pragma unroll
      for (int n = 0; n < 2; ++n) s[ti][n][0] = s[ti][n][1] = s[ti][n][2] = s[ti][n][3] = 0.f;
      if (ti < nt) {
        const uint32_t kb = smem_u32(sm.ring[slot][ti]);
        if (FP8) {
#pragma unroll
          for (int c = 0; c < 8; c += 2) {  // 16-byte chunks c, c+1 = k-steps c, c+1
            uint32_t r0, r1, r2, r3;  // (tok 0-7, c) (tok 0-7, c+1) (tok 8-15, c) (tok 8-15, c+1)
            ldsm_x4(kb + pl.k[c >> 1], r0, r1, r2, r3);
            mma_f16_16816(s[ti][0], qa[c][0], qa[c][1], qa[c][2], qa[c][3], e4m3_lo_f16x2(r0), e4m3_hi_f16x2(r0));
            mma_f16_16816(s[ti][0], qa[c + 1][0], qa[c + 1][1], qa[c + 1][2], qa[c + 1][3], e4m3_lo_f16x2(r1),
                          e4m3_hi_f16x2(r1));
            mma_f16_16816(s[ti][1], qa[c][0], qa[c][1], qa[c][2], qa[c][3], e4m3_lo_f16x2(r2), e4m3_hi_f16x2(r2));
            mma_f16_16816(s[ti][1], qa[c + 1][0], qa[c + 1][1], qa[c + 1][2], qa[c + 1][3], e4m3_lo_f16x2(r3),
                          e4m3_hi_f16x2(r3));
          }
        } else {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            uint32_t r0, r1, r2, r3;  // b0,b1 of tokens 0-7; b0,b1 of tokens 8-15
            ldsm_x4(kb + (kk >> 2) * kHalfBytes + pl.k[kk & 3], r0, r1, r2, r3);
            mma_bf16_16816(s[ti][0], qa[kk][0], qa[kk][1], qa[kk][2], qa[kk][3], r0, r1);
            mma_bf16_16816(s[ti][1], qa[kk][0], qa[kk][1], qa[kk][2], qa[kk][3], r2, r3);
          }
        }
      }
    }
    // ---- scale, causal mask (only the stages that reach the diagonal or the sequence end)
    const int j_base = st * kPfKt * kTileTokens;
    const bool need_mask = (j_base + kPfKt * kTileTokens - 1 > ctx0 + i0) || nt < kPfKt;
#pragma unroll
    for (int ti = 0; ti < kPfKt; ++ti)
#pragma unroll
      for (int n = 0; n < 2; ++n) {
        const int j = j_base + ti * kTileTokens + n * 8 + 2 * t;
        float* c = s[ti][n];
        c[0] *= scale, c[1] *= scale, c[2] *= scale, c[3] *= scale;
        if (need_mask) {
          const bool dead = ti >= nt;
          if (dead || j > pos0) c[0] = -CUDART_INF_F;
          if (dead || j + 1 > pos0) c[1] = -CUDART_INF_F;
          if (dead || j > pos1) c[2] = -CUDART_INF_F;
          if (dead || j + 1 > pos1) c[3] = -CUDART_INF_F;
        }
      }
    // ---- online softmax with a lazily refreshed reference (see kTau)
    float mx0 = -CUDART_INF_F, mx1 = -CUDART_INF_F;
#pragma unroll
    for (int ti = 0; ti < kPfKt; ++ti)
#pragma unroll
      for (int n = 0; n < 2; ++n) {
        mx0 = fmaxf(mx0, fmaxf(s[ti][n][0], s[ti][n][1]));
        mx1 = fmaxf(mx1, fmaxf(s[ti][n][2], s[ti][n][3]));
      }
    const bool grow = (mx0 > m0 + kTau) | (mx1 > m1 + kTau);
    if (__any_sync(0xffffffffu, grow)) {
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
      const float al0 = (mn0 == -CUDART_INF_F) ? 1.f : fast_exp2(m0 - mn0);
      const float al1 = (mn1 == -CUDART_INF_F) ? 1.f : fast_exp2(m1 - mn1);
      m0 = mn0, m1 = mn1;
      l0 *= al0, l1 *= al1;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        o[j][0] *= al0, o[j][1] *= al0;
        o[j][2] *= al1, o[j][3] *= al1;
      }
    }
    const float mu0 = (m0 == -CUDART_INF_F) ? 0.f : m0, mu1 = (m1 == -CUDART_INF_F) ? 0.f : m1;
    // ---- O += P V per tile
#pragma unroll
    for (int ti = 0; ti < kPfKt; ++ti) {
      if (ti < nt) {
        float* a = s[ti][0];
        float* c = s[ti][1];
        const float p00 = fast_exp2(a[0] - mu0), p01 = fast_exp2(a[1] - mu0);
        const float p02 = fast_exp2(a[2] - mu1), p03 = fast_exp2(a[3] - mu1);
        const float p10 = fast_exp2(c[0] - mu0), p11 = fast_exp2(c[1] - mu0);
        const float p12 = fast_exp2(c[2] - mu1), p13 = fast_exp2(c[3] - mu1);
        l0 += p00 + p01 + p10 + p11;
        l1 += p02 + p03 + p12 + p13;
        uint32_t pa0, pa1, pa2, pa3;  // P as A operand: rows g / g+8, kv tokens 2t,+1 / 2t+8,+9
        if (FP8) {
          pa0 = pack_f16x2(p00, p01), pa1 = pack_f16x2(p02, p03);
          pa2 = pack_f16x2(p10, p11), pa3 = pack_f16x2(p12, p13);
        } else {
          pa0 = pack_bf16x2(p00, p01), pa1 = pack_bf16x2(p02, p03);
          pa2 = pack_bf16x2(p10, p11), pa3 = pack_bf16x2(p12, p13);
        }
        // tokens of this tile beyond the sequence end hold whatever the page holds: keep them out
        // of the accumulation even as 0 x NaN by clearing their halves of the V fragments
        const int valid = kv_end - (j_base + ti * kTileTokens);  // > 0
        const uint32_t vb = smem_u32(sm.ring[slot][ti]) + TB;
        if (FP8) {
#pragma unroll
          for (int jp = 0; jp < 4; ++jp) {  // chunks 2jp, 2jp+1 -> d-tiles 4jp .. 4jp+3
            uint32_t r0, r1, r2, r3;  // (tok 0-7, 2jp) (tok 8-15, 2jp) (tok 0-7, 2jp+1) (tok 8-15, 2jp+1)
            ldsm_x4_trans(vb + pl.v[jp], r0, r1, r2, r3);
            if (valid < kTileTokens) {
              const uint32_t k0 = (2 * t < valid ? 0x0000ffffu : 0u) | (2 * t + 1 < valid ? 0xffff0000u : 0u);
              const uint32_t k1 = (2 * t + 8 < valid ? 0x0000ffffu : 0u) | (2 * t + 9 < valid ? 0xffff0000u : 0u);
              r0 &= k0, r2 &= k0, r1 &= k1, r3 &= k1;
            }
            mma_f16_16816(o[4 * jp], pa0, pa1, pa2, pa3, e4m3_lo_f16x2(__byte_perm(r0, 0, 0x5420)),
                          e4m3_lo_f16x2(__byte_perm(r1, 0, 0x5420)));
            mma_f16_16816(o[4 * jp + 1], pa0, pa1, pa2, pa3, e4m3_lo_f16x2(__byte_perm(r0, 0, 0x5431)),
                          e4m3_lo_f16x2(__byte_perm(r1, 0, 0x5431)));
            mma_f16_16816(o[4 * jp + 2], pa0, pa1, pa2, pa3, e4m3_lo_f16x2(__byte_perm(r2, 0, 0x5420)),
                          e4m3_lo_f16x2(__byte_perm(r3, 0, 0x5420)));
            mma_f16_16816(o[4 * jp + 3], pa0, pa1, pa2, pa3, e4m3_lo_f16x2(__byte_perm(r2, 0, 0x5431)),
                          e4m3_lo_f16x2(__byte_perm(r3, 0, 0x5431)));
          }
        } else {
#pragma unroll
          for (int jp = 0; jp < 8; ++jp) {  // d-tiles 2jp, 2jp+1
            uint32_t r0, r1, r2, r3;  // b0,b1 of d-tile 2jp; b0,b1 of d-tile 2jp+1
            ldsm_x4_trans(vb + (jp >> 2) * kHalfBytes + pl.v[jp & 3], r0, r1, r2, r3);
            if (valid < kTileTokens) {
              const uint32_t k0 = (2 * t < valid ? 0x0000ffffu : 0u) | (2 * t + 1 < valid ? 0xffff0000u : 0u);
              const uint32_t k1 = (2 * t + 8 < valid ? 0x0000ffffu : 0u) | (2 * t + 9 < valid ? 0xffff0000u : 0u);
              r0 &= k0, r2 &= k0, r1 &= k1, r3 &= k1;
            }
            mma_bf16_16816(o[2 * jp], pa0, pa1, pa2, pa3, r0, r1);
            mma_bf16_16816(o[2 * jp + 1], pa0, pa1, pa2, pa3, r2, r3);
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.empty[slot]);
  }

  // ---- epilogue: normalise and store the two rows of this thread
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  float inv0 = l0 > 0.f ? 1.f / l0 : 0.f, inv1 = l1 > 0.f ? 1.f / l1 : 0.f;
  if (FP8) {
    const float vs = __ldg(p.v_scale + kvh * p.scale_stride);
    inv0 *= vs, inv1 *= vs;
  }
  __nv_bfloat16* d0 = p.out + static_cast<int64_t>(q0 + i0 + g) * p.out_stride_tok + static_cast<int64_t>(h) * kHeadDim;
  __nv_bfloat16* d1 = d0 + 8 * p.out_stride_tok;
  if (FP8) {
    // d-tile 2c holds d = 16c + 2n, d-tile 2c+1 holds d = 16c + 2n + 1 (n = 2t, 2t+1): 4 consecutive d
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      if (row0_ok) {
        uint2 pk;
        pk.x = pack_bf16x2(o[2 * c][0] * inv0, o[2 * c + 1][0] * inv0);
        pk.y = pack_bf16x2(o[2 * c][1] * inv0, o[2 * c + 1][1] * inv0);
        *reinterpret_cast<uint2*>(d0 + 16 * c + 4 * t) = pk;
      }
      if (row1_ok) {
        uint2 pk;
        pk.x = pack_bf16x2(o[2 * c][2] * inv1, o[2 * c + 1][2] * inv1);
        pk.y = pack_bf16x2(o[2 * c][3] * inv1, o[2 * c + 1][3] * inv1);
        *reinterpret_cast<uint2*>(d1 + 16 * c + 4 * t) = pk;
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (row0_ok) *reinterpret_cast<uint32_t*>(d0 + 8 * j + 2 * t) = pack_bf16x2(o[j][0] * inv0, o[j][1] * inv0);
      if (row1_ok) *reinterpret_cast<uint32_t*>(d1 + 8 * j + 2 * t) = pack_bf16x2(o[j][2] * inv1, o[j][3] * inv1);
    }
  }
}

size_t prefill_smem_bytes(bool fp8) {
  return (fp8 ? sizeof(PfSmem<2 * kTileBytesFp8>) : sizeof(PfSmem<2 * kTileBytesBf16>)) + 1024;
}

template <typename K>
static cudaError_t pf_launch(K kern, int* configured_dev, size_t smem, const CUtensorMap& tm_k, const CUtensorMap& tm_v,
                             const PrefillParams& p, dim3 grid, cudaStream_t stream) {
  int dev = 0;
  if (cudaError_t e = cudaGetDevice(&dev)) return e;
  if (dev != *configured_dev) {
    if (cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)))
      return e;
    *configured_dev = dev;
  }
  kern<<<grid, (kPfWarps + 1) * 32, smem, stream>>>(tm_k, tm_v, p);
  return cudaGetLastError();
}

cudaError_t launch_paged_prefill(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const PrefillParams& p, bool fp8,
                                 cudaStream_t stream) {
  static thread_local int dev_bf16 = -1, dev_fp8 = -1;
  // tile slots: floor(num_tokens / 16) + batch >= sum of ceil(q_len / 16) (see pf_locate)
  const dim3 grid(static_cast<unsigned>(p.num_tokens / kPfTok + p.batch),
                  static_cast<unsigned>(p.num_kv_heads * p.subgroups));
  if (fp8)
    return pf_launch(paged_prefill_kernel<true>, &dev_fp8, prefill_smem_bytes(true), tm_k, tm_v, p, grid, stream);
  return pf_launch(paged_prefill_kernel<false>, &dev_bf16, prefill_smem_bytes(false), tm_k, tm_v, p, grid, stream);
}

}  // namespace b200
