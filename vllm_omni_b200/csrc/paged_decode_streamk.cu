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
    zero_v_tail<1>(stage, valid, lane);
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
    sk_zero_v_tail_fp8(stage, valid, lane);
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
        sk_consume_fp8(ring0 + cslot * SB, sm.ring[warp][cslot], valid, qf, sl, scale, st, lane);
      else
        sk_consume(ring0 + cslot * SB, sm.ring[warp][cslot], valid, qf, sl, scale, st, lane);
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
