// Per-warp tile math shared by the stream-K decode kernel and the paged prefill kernel: one
// 16-token KV tile (SWIZZLE_128B rows in shared memory) against 8 query columns held as MMA
// B-fragments, with a lazily refreshed online-softmax reference, for bf16 and fp8-e4m3 caches,
// and the epilogue that normalises and writes a finished (or partial) row.
#pragma once
#include "decode_common.cuh"

namespace b200 {

// Per-lane shared-memory byte offsets inside a stage, computed once.
struct SkLane {
  uint32_t kq[4];  // ldmatrix of K (A operand), k-step mod 4
  uint32_t vq[4];  // ldmatrix.trans of V, d-tile mod 4 (relative to the V tile)
};
__device__ __forceinline__ SkLane sk_make_lane(int lane) {
  SkLane o;
  const LaneOffsets lo = make_lane_offsets(lane);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    o.kq[i] = lo.k[i];
    o.vq[i] = lo.v[i] + kTileBytesBf16;
  }
  return o;
}

__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2,
                                            int c3, int c4, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      ".L2::cache_hint [%0], [%1, {%3, %4, %5, %6, %7}], [%2], %8;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "l"(policy)
      : "memory");
}

__device__ __forceinline__ void tma_load_4d_s(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2,
                                              int c3, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      ".L2::cache_hint [%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(policy)
      : "memory");
}

// Online-softmax reference max is refreshed lazily: the running reference m may lag the true
// row max by up to kTau (log2 units), so P <= 2^kTau (exact in fp32 / bf16 range) and the
// accumulator rescale — and the cross-lane max reduction that feeds it — only run when some
// score outgrows the reference. The result is the same softmax (any reference cancels).
constexpr float kTau = 8.f;

// S^T = K Q^T, softmax update, O^T += V^T P^T for one 16-token tile in the ring slot at `base`.
// ZERO_TAIL: the caller owns the stage exclusively, so the V rows behind `valid` may be zeroed in
// place (they can hold non-finite garbage when they lie behind the end of the sequence); a caller
// that shares the stage between warps zeroes the rows behind the sequence end itself.
template <bool ZERO_TAIL>
__device__ __forceinline__ void sk_consume(uint32_t base, uint8_t* stage, int valid, const uint32_t (&qf)[8][2],
                                           const SkLane& sl, float scale_log2, WarpState& st, int lane) {
  float sa[4] = {0.f, 0.f, 0.f, 0.f}, sb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int kk = 0; kk < 8; kk += 2) {  // two independent accumulation chains
    uint32_t a0, a1, a2, a3, b0, b1, b2, b3;
    ldsm_x4(base + (kk >> 2) * kHalfBytes + sl.kq[kk & 3], a0, a1, a2, a3);
    ldsm_x4(base + (kk >> 2) * kHalfBytes + sl.kq[(kk + 1) & 3], b0, b1, b2, b3);
    mma_bf16_16816(sa, a0, a1, a2, a3, qf[kk][0], qf[kk][1]);
    mma_bf16_16816(sb, b0, b1, b2, b3, qf[kk + 1][0], qf[kk + 1][1]);
  }
  // scaled scores (log2 domain): rows g (x0,x1: heads 2t,2t+1) and g+8 (x2,x3)
  float x0 = (sa[0] + sb[0]) * scale_log2, x1 = (sa[1] + sb[1]) * scale_log2;
  float x2 = (sa[2] + sb[2]) * scale_log2, x3 = (sa[3] + sb[3]) * scale_log2;
  if (valid < kTileTokens) {  // last tile of a sequence: mask the tail, keep V finite
    const int g = lane >> 2;
    if (g >= valid) x0 = x1 = -CUDART_INF_F;
    if (g + 8 >= valid) x2 = x3 = -CUDART_INF_F;
    if (ZERO_TAIL) zero_v_tail<1>(stage, valid, lane);
  }
  const bool grow = (fmaxf(x0, x2) > st.m[0] + kTau) | (fmaxf(x1, x3) > st.m[1] + kTau);
  if (__any_sync(0xffffffffu, grow)) {
    float mx0 = fmaxf(x0, x2), mx1 = fmaxf(x1, x3);
#pragma unroll
    for (int off = 4; off <= 16; off <<= 1) {
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, off));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, off));
    }
    const float mn0 = fmaxf(st.m[0], mx0), mn1 = fmaxf(st.m[1], mx1);
    const float al0 = (mn0 == -CUDART_INF_F) ? 1.f : fast_exp2(st.m[0] - mn0);
    const float al1 = (mn1 == -CUDART_INF_F) ? 1.f : fast_exp2(st.m[1] - mn1);
    st.m[0] = mn0;
    st.m[1] = mn1;
    st.l[0] *= al0;
    st.l[1] *= al1;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      st.acc[j][0] *= al0;
      st.acc[j][1] *= al1;
      st.acc[j][2] *= al0;
      st.acc[j][3] *= al1;
    }
  }
  const float mu0 = (st.m[0] == -CUDART_INF_F) ? 0.f : st.m[0];
  const float mu1 = (st.m[1] == -CUDART_INF_F) ? 0.f : st.m[1];
  const float p0 = fast_exp2(x0 - mu0), p1 = fast_exp2(x1 - mu1);
  const float p2 = fast_exp2(x2 - mu0), p3 = fast_exp2(x3 - mu1);
  st.l[0] += p0 + p2;
  st.l[1] += p1 + p3;
  const uint32_t pb0 = movmatrix_trans(pack_bf16x2(p0, p1));  // tokens 0-7  -> B-frag k 0-7
  const uint32_t pb1 = movmatrix_trans(pack_bf16x2(p2, p3));  // tokens 8-15 -> B-frag k 8-15
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint32_t a0, a1, a2, a3;
    ldsm_x4_trans(base + (j >> 2) * kHalfBytes + sl.vq[j & 3], a0, a1, a2, a3);
    mma_bf16_16816(st.acc[j], a0, a1, a2, a3, pb0, pb1);
  }
}

// ------------------------------------------------------------------------------------------------
// FP8-e4m3 KV cache (per-tensor or per-kv-head scales, vllm_omni/quantization -> vLLM
// BaseKVCacheMethod / compressed-tensors attn_head strategy): a stage is a 16-token K tile and
// V tile of 128 B rows (2 KiB each, SWIZZLE_128B). e4m3 -> f16 is exact, so the math runs as
// f16 tensor-core MMAs on register-converted fragments; Q is held in f16 (bf16's 8 significant
// bits fit), the K scale is folded into the softmax scale and the V scale into the output.
//
// Fragment bookkeeping (b16 = two adjacent e4m3 d-values):
//  * K (A operand, rows = tokens): one ldmatrix.x4 returns, for lane (g, t), the 4 bytes
//    d = 16c + 4t .. 4t+3 of token g (r0), token g+8 (r1) for chunk c, and the same for chunk
//    c+1 (r2, r3). The contraction index inside a 16-wide k-step is permuted accordingly:
//    MMA k = (2t, 2t+1) <- d = 16c + 4t + (0, 1), k = (2t+8, 2t+9) <- d = 16c + 4t + (2, 3);
//    Q's B fragments are loaded with the same permutation (sk_load_q_f16).
//  * V (A operand of O^T = V^T P^T, rows = d): ldmatrix.x4.trans returns bytes
//    (tok 2t, d 2g), (tok 2t, d 2g+1), (tok 2t+1, d 2g), (tok 2t+1, d 2g+1); a byte permute
//    regroups them into token pairs per d. MMA row g <- d = 16j + 2g, row g+8 <- d = 16j + 2g + 1
//    (sk_store undoes this row permutation).
// ------------------------------------------------------------------------------------------------
constexpr int kTileBytesFp8 = kTileTokens * kHeadDim;  // 2048: one K (or V) tile
constexpr int kKvTileBytesFp8 = 2 * kTileBytesFp8;

__device__ __forceinline__ uint32_t e4m3_lo_f16x2(uint32_t r) {
  return cvt_e4m3x2_to_f16x2(static_cast<uint16_t>(r & 0xffffu));
}
__device__ __forceinline__ uint32_t e4m3_hi_f16x2(uint32_t r) {
  return cvt_e4m3x2_to_f16x2(static_cast<uint16_t>(r >> 16));
}

// Q^T B-fragments in f16 with the fp8 k-permutation: qf[kk][0] = Q[head g][16kk + 4t + (0,1)],
// qf[kk][1] = Q[head g][16kk + 4t + (2,3)]; heads beyond the group are zero.
__device__ __forceinline__ void sk_load_q_f16(const DecodeParams& p, int b, int kvh, int sub, int lane,
                                              uint32_t (&qf)[8][2]) {
  const int g = lane >> 2, t = lane & 3;
  const int head_in_group = sub * kHeadsPerItem + g;
  const bool valid = head_in_group < p.group;
  const int h = kvh * p.group + head_in_group;
  const uint32_t* row = reinterpret_cast<const uint32_t*>(p.q + static_cast<int64_t>(b) * p.q_stride_tok +
                                                          static_cast<int64_t>(h) * kHeadDim);
#pragma unroll
  for (int kk = 0; kk < 8; ++kk) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const uint32_t raw = valid ? __ldg(row + kk * 8 + 2 * t + i) : 0u;  // two bf16
      const float lo = __uint_as_float(raw << 16), hi = __uint_as_float(raw & 0xffff0000u);
      qf[kk][i] = pack_f16x2(lo, hi);
    }
  }
}

// Zero V rows [valid, 16) of an fp8 stage (a whole 128 B row each, so the swizzle is irrelevant).
__device__ __forceinline__ void sk_zero_v_tail_fp8(uint8_t* stage, int valid, int lane) {
  for (int idx = lane; idx < kTileTokens * 8; idx += 32) {
    const int row = idx >> 3;
    if (row >= valid)
      *reinterpret_cast<int4*>(stage + kTileBytesFp8 + row * 128 + (idx & 7) * 16) = make_int4(0, 0, 0, 0);
  }
  __syncwarp();
}

template <bool ZERO_TAIL>
__device__ __forceinline__ void sk_consume_fp8(uint32_t base, uint8_t* stage, int valid, const uint32_t (&qf)[8][2],
                                               const SkLane& sl, float scale_log2, WarpState& st, int lane) {
  float sa[4] = {0.f, 0.f, 0.f, 0.f}, sb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    uint32_t r0, r1, r2, r3;
    ldsm_x4(base + sl.kq[c >> 1], r0, r1, r2, r3);
    mma_f16_16816(sa, e4m3_lo_f16x2(r0), e4m3_lo_f16x2(r1), e4m3_hi_f16x2(r0), e4m3_hi_f16x2(r1), qf[c][0], qf[c][1]);
    mma_f16_16816(sb, e4m3_lo_f16x2(r2), e4m3_lo_f16x2(r3), e4m3_hi_f16x2(r2), e4m3_hi_f16x2(r3), qf[c + 1][0],
                  qf[c + 1][1]);
  }
  float x0 = (sa[0] + sb[0]) * scale_log2, x1 = (sa[1] + sb[1]) * scale_log2;
  float x2 = (sa[2] + sb[2]) * scale_log2, x3 = (sa[3] + sb[3]) * scale_log2;
  if (valid < kTileTokens) {
    const int g = lane >> 2;
    if (g >= valid) x0 = x1 = -CUDART_INF_F;
    if (g + 8 >= valid) x2 = x3 = -CUDART_INF_F;
    if (ZERO_TAIL) sk_zero_v_tail_fp8(stage, valid, lane);
  }
  const bool grow = (fmaxf(x0, x2) > st.m[0] + kTau) | (fmaxf(x1, x3) > st.m[1] + kTau);
  if (__any_sync(0xffffffffu, grow)) {
    float mx0 = fmaxf(x0, x2), mx1 = fmaxf(x1, x3);
#pragma unroll
    for (int off = 4; off <= 16; off <<= 1) {
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, off));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, off));
    }
    const float mn0 = fmaxf(st.m[0], mx0), mn1 = fmaxf(st.m[1], mx1);
    const float al0 = (mn0 == -CUDART_INF_F) ? 1.f : fast_exp2(st.m[0] - mn0);
    const float al1 = (mn1 == -CUDART_INF_F) ? 1.f : fast_exp2(st.m[1] - mn1);
    st.m[0] = mn0;
    st.m[1] = mn1;
    st.l[0] *= al0;
    st.l[1] *= al1;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      st.acc[j][0] *= al0;
      st.acc[j][1] *= al1;
      st.acc[j][2] *= al0;
      st.acc[j][3] *= al1;
    }
  }
  const float mu0 = (st.m[0] == -CUDART_INF_F) ? 0.f : st.m[0];
  const float mu1 = (st.m[1] == -CUDART_INF_F) ? 0.f : st.m[1];
  const float p0 = fast_exp2(x0 - mu0), p1 = fast_exp2(x1 - mu1);
  const float p2 = fast_exp2(x2 - mu0), p3 = fast_exp2(x3 - mu1);
  st.l[0] += p0 + p2;
  st.l[1] += p1 + p3;
  const uint32_t pb0 = movmatrix_trans(pack_f16x2(p0, p1));
  const uint32_t pb1 = movmatrix_trans(pack_f16x2(p2, p3));
#pragma unroll
  for (int jp = 0; jp < 4; ++jp) {
    uint32_t r0, r1, r2, r3;  // (tok 0-7, chunk 2jp) (tok 8-15, 2jp) (tok 0-7, 2jp+1) (tok 8-15, 2jp+1)
    ldsm_x4_trans(base + kTileBytesFp8 + sl.kq[jp], r0, r1, r2, r3);
    mma_f16_16816(st.acc[2 * jp], e4m3_lo_f16x2(__byte_perm(r0, 0, 0x5420)), e4m3_lo_f16x2(__byte_perm(r0, 0, 0x5431)),
                  e4m3_lo_f16x2(__byte_perm(r1, 0, 0x5420)), e4m3_lo_f16x2(__byte_perm(r1, 0, 0x5431)), pb0, pb1);
    mma_f16_16816(st.acc[2 * jp + 1], e4m3_lo_f16x2(__byte_perm(r2, 0, 0x5420)),
                  e4m3_lo_f16x2(__byte_perm(r2, 0, 0x5431)), e4m3_lo_f16x2(__byte_perm(r3, 0, 0x5420)),
                  e4m3_lo_f16x2(__byte_perm(r3, 0, 0x5431)), pb0, pb1);
  }
}

// Normalise the warp's state of one item segment and write it: the final bf16 row (+ LSE) when
// the segment is the whole item, else a partial into workspace slot `piece`. `stage` is a free
// ring slot of this warp used to transpose the fragments into rows.
template <bool FP8>
__device__ __forceinline__ void sk_store(const DecodeParams& p, WarpState& st, int b, int hs, bool whole, int piece,
                                         float* stage, int lane) {
  const int g = lane >> 2, t = lane & 3;
#pragma unroll
  for (int off = 4; off <= 16; off <<= 1) {
    st.l[0] += __shfl_xor_sync(0xffffffffu, st.l[0], off);
    st.l[1] += __shfl_xor_sync(0xffffffffu, st.l[1], off);
  }
  float inv0 = st.l[0] > 0.f ? 1.f / st.l[0] : 0.f;
  float inv1 = st.l[1] > 0.f ? 1.f / st.l[1] : 0.f;
  if (FP8) {  // dequantisation scale of V (value = e4m3 * v_scale[kv head])
    const float vs = __ldg(p.v_scale + (hs / p.subgroups) * p.scale_stride);
    inv0 *= vs;
    inv1 *= vs;
  }
  // fragment row -> d: bf16 tiles (g, g+8); fp8 tiles (2g, 2g+1), see sk_consume_fp8
  const int da = FP8 ? 2 * g : g, db = FP8 ? 2 * g + 1 : g + 8;
  const float lse0 = st.l[0] > 0.f ? (st.m[0] + log2f(st.l[0])) * kLn2 : -CUDART_INF_F;
  const float lse1 = st.l[1] > 0.f ? (st.m[1] + log2f(st.l[1])) * kLn2 : -CUDART_INF_F;
  const int kvh = hs / p.subgroups, sub = hs % p.subgroups;
  // two passes of four heads: the staging rows (4 x 132 floats + 4 LSE) fit the smallest (fp8,
  // 4 KiB) ring slot
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    if ((t >> 1) == half) {
      const int r0 = (2 * t) & 3, r1 = r0 + 1;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        stage[r0 * kMergeRow + 16 * j + da] = st.acc[j][0] * inv0;
        stage[r1 * kMergeRow + 16 * j + da] = st.acc[j][1] * inv1;
        stage[r0 * kMergeRow + 16 * j + db] = st.acc[j][2] * inv0;
        stage[r1 * kMergeRow + 16 * j + db] = st.acc[j][3] * inv1;
      }
      if (g == 0) {  // natural-log LSE behind the rows
        stage[4 * kMergeRow + r0] = lse0;
        stage[4 * kMergeRow + r1] = lse1;
      }
    }
    __syncwarp();
#pragma unroll
    for (int hh = 0; hh < 4; ++hh) {
      const int head_in_group = sub * kHeadsPerItem + half * 4 + hh;
      if (head_in_group >= p.group) break;
      const int h = kvh * p.group + head_in_group;
      const float4 v = *reinterpret_cast<const float4*>(stage + hh * kMergeRow + lane * 4);
      const float lse = stage[4 * kMergeRow + hh];
      if (whole) {
        uint2 pk;
        pk.x = pack_bf16x2(v.x, v.y);
        pk.y = pack_bf16x2(v.z, v.w);
        *reinterpret_cast<uint2*>(p.out + static_cast<int64_t>(b) * p.out_stride_tok +
                                  static_cast<int64_t>(h) * kHeadDim + lane * 4) = pk;
        if (p.lse != nullptr && lane == 0) p.lse[static_cast<int64_t>(b) * p.num_q_heads + h] = lse;
      } else {
        const int64_t row = static_cast<int64_t>(piece) * kHeadsPerItem + half * 4 + hh;
        *reinterpret_cast<float4*>(p.part_o + row * kHeadDim + lane * 4) = v;
        if (lane == 0) p.part_lse[row] = lse;
      }
    }
    __syncwarp();
  }
}

}  // namespace b200
