// Causal varlen attention over the paged KV cache on the 5th-generation tensor cores (tcgen05 / UMMA,
// accumulators in TMEM) — the "dense enough to pay off" case of GQA query-group packing: 16 query tokens x the
// 8 query heads of one KV head = 128 rows are exactly one M = 128 UMMA tile with no padding at G = 8, and a CTA
// owns TWO such tiles (32 consecutive query tokens) that share every staged K/V block, which halves the L2 -> SM
// traffic that bounded the one-tile version (measured 408 TFLOP/s at 3.1 TB/s of L2 reads).
// FP8 caches reach this kernel through the bf16 staging of cache_ops.cu (stage_fp8_pages_kernel, scales applied
// here); the mma.sync kernel in paged_prefill.cu is the unstaged fp8 path and the comparison point.
// Same call as paged_prefill.cu (vllm/v1/attention/backends/flash_attn.py:758-819 with max_query_len > 1).
//
// Per KV block of 64 tokens (4 pages of 16) and per tile t:
//   S_t[128 x 64]  = Q_t[128 x 128d] . K^T   8 x tcgen05.mma m128 n64 k16, A and B K-major from shared memory
//   P_t = exp2(S_t * scale - m) (causal / tail masked), row statistics in registers: thread r of warpgroup t owns
//         row r (TMEM lane r)
//   O_t[128 x 128d] += P_t[128 x 64] . V     4 x tcgen05.mma m128 n128 k16, A = P_t read from TMEM (the softmax threads
//                                            store it as bf16 pairs over the first 32 columns of the S buffer they
//                                            just consumed), B = V MN-major from shared memory, both 64-wide d
//                                            halves in one instruction through the descriptor's leading byte offset
// K/V pages are staged by TMA straight into the SWIZZLE_128B layout the UMMA descriptors describe (two boxes of
// 64 d x 16 tokens per page, one per 128-byte half of the head dimension), four stages of 32 KiB (a stage is held
// from its load until P.V of its block completes, so four are needed to give a load two block-times to land). S is
// double-buffered in TMEM (2 tiles x 2 x 64 columns), so Q.K^T of block j+1 and P.V of block j-1
// run under the softmax of block j and the two tiles' softmax warpgroups alternate on the tensor core; O lives in
// 2 x 128 more columns (512 in all) and is rescaled in TMEM only when a row maximum grows by more than 2^kTau.
//
// Warp roles (320 threads): warps 0-3 / 4-7 softmax, correction and epilogue of tile 0 / 1 (warp w touches TMEM
// lanes 32 (w % 4) ..), warp 8 TMA producer, warp 9 TMEM allocation + single-thread MMA issue.
#include "kernels.h"
#include "sk_math.cuh"
#include "umma.cuh"

#ifndef B200_UMMA_P_IN_TMEM
#define B200_UMMA_P_IN_TMEM 1  // P as the TMEM A operand of P.V (0: through shared memory)
#endif

namespace b200 {

constexpr int kUpQT = 2;               // 128-row query tiles per CTA
constexpr int kUpTok = 16;             // query tokens per tile
constexpr int kUpCtaTok = kUpQT * kUpTok;
constexpr int kUpHeads = 8;            // query-head slots per tile (rows = token * 8 + head slot)
constexpr int kUpBlk = 64;             // KV tokens per block
constexpr int kUpTiles = kUpBlk / kTileTokens;  // 16-token pages per block
constexpr int kUpQHalf = 128 * 128;    // Q / P operand half: [128 rows][128 B] = 16 KiB
constexpr int kUpKvHalf = kUpBlk * 128;  // K / V operand half: [64 tokens][128 B] = 8 KiB
constexpr int kUpStages = 4;
constexpr int kUpThreads = kUpQT * 128 + 64;
constexpr uint32_t kUpTmemCols = 512;  // S[t][buf]: 64 columns at t * 128 + buf * 64; O[t]: 128 columns at 256 + t * 128

struct UpSmem {
  alignas(1024) uint8_t q[kUpQT][2][kUpQHalf];     // A of Q.K^T: [tile][d half][row][64 d]
#if !B200_UMMA_P_IN_TMEM
  alignas(1024) uint8_t p[kUpQT][kUpQHalf];        // A of P.V: [tile][row][64 tokens]
#endif
  alignas(1024) uint8_t k[kUpStages][2][kUpKvHalf];  // B of Q.K^T: [d half][token][64 d]
  alignas(1024) uint8_t v[kUpStages][2][kUpKvHalf];  // B of P.V (MN-major): [d half][token][64 d]
  alignas(8) uint64_t kv_full[kUpStages];
  alignas(8) uint64_t kv_empty[kUpStages];
  alignas(8) uint64_t s_full[kUpQT][2];
  alignas(8) uint64_t p_full[kUpQT];
  alignas(8) uint64_t pv_done[kUpQT];  // P_t.V of block j has completed (waited for once per block, in order)
  uint32_t tmem_base;
};

__device__ __forceinline__ void tma_load_5d_half(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1,
                                                 int c2, int c3, int c4, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      ".L2::cache_hint [%0], [%1, {%3, %4, %5, %6, %7}], [%2], %8;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "l"(policy)
      : "memory");
}

// CTA -> (sequence, 32-token query chunk), the slot scheme of paged_prefill.cu (pf_locate) at 32-token granularity:
// sequence b owns slots [qsl[b] / 32 + b, qsl[b+1] / 32 + b + 1), always enough for its ceil(q_len / 32) chunks.
__device__ __forceinline__ bool up_locate(const PrefillParams& p, int slot, int& b, int& qt, int& q0, int& q_len) {
  int lo = 0, hi = p.batch;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if ((__ldg(p.query_start_loc + mid) >> 5) + mid <= slot) lo = mid; else hi = mid;
  }
  b = lo;
  q0 = __ldg(p.query_start_loc + b);
  q_len = __ldg(p.query_start_loc + b + 1) - q0;
  qt = slot - ((q0 >> 5) + b);
  return qt * kUpCtaTok < q_len;
}

__global__ void __launch_bounds__(kUpThreads, 1)
paged_prefill_umma_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                          const PrefillParams p) {
  extern __shared__ uint8_t smem_raw[];
  UpSmem& sm = *reinterpret_cast<UpSmem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kProducerWarp = kUpQT * 4, kMmaWarp = kUpQT * 4 + 1;

  int b, qt, q0, q_len;
  // heavy (late) chunks first; the whole CTA leaves together, before any barrier or TMEM allocation
  if (!up_locate(p, static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x), b, qt, q0, q_len)) return;
  const int hs = blockIdx.y;
  const int kvh = hs / p.subgroups, sub = hs % p.subgroups;
  const int heads_here = min(kUpHeads, p.group - sub * kUpHeads);
  const int S = max(0, __ldg(p.seq_lens + b));
  const int i0 = qt * kUpCtaTok;
  const int ctx0 = S - q_len;  // cache position of the sequence's first new token
  const int kv_end = max(0, min(S, ctx0 + min(q_len, i0 + kUpCtaTok)));  // the chunk's last query sees [0, kv_end)
  const int ntiles = (kv_end + kTileTokens - 1) / kTileTokens;
  const int nblk = (kv_end + kUpBlk - 1) / kUpBlk;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kUpStages; ++s) {
      mbar_init(&sm.kv_full[s], 1);
      mbar_init(&sm.kv_empty[s], 1);
    }
#pragma unroll
    for (int t = 0; t < kUpQT; ++t) {
      mbar_init(&sm.s_full[t][0], 1);
      mbar_init(&sm.s_full[t][1], 1);
      mbar_init(&sm.p_full[t], 128);
      mbar_init(&sm.pv_done[t], 1);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
  }
  if (warp == kMmaWarp) tmem_alloc<kUpTmemCols>(&sm.tmem_base);
  // ---- Q tiles -> shared memory (swizzled K-major A operand); rows of absent tokens / head slots are zero
  if (warp < kUpQT * 4) {
    const int t = warp >> 2, r = threadIdx.x & 127, i = t * kUpTok + (r >> 3), hh = r & 7;
    const bool ok = (i0 + i < q_len) && (hh < heads_here);
    const int h = kvh * p.group + sub * kUpHeads + hh;
    const int4* src = reinterpret_cast<const int4*>(p.q + static_cast<int64_t>(q0 + i0 + i) * p.q_stride_tok +
                                                    static_cast<int64_t>(h) * kHeadDim);
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      const int4 x = ok ? __ldg(src + c) : make_int4(0, 0, 0, 0);
      *reinterpret_cast<int4*>(&sm.q[t][c >> 3][r * 128 + (((c & 7) ^ (r & 7)) << 4)]) = x;
    }
    fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's async-proxy reads
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = sm.tmem_base;

  if (warp == kProducerWarp) {
    // ===================== TMA producer =====================
    const uint64_t pol = l2_policy_evict_last();  // the other query chunks of the sequence re-read these pages
    int page0 = -32, ids = 0;
    for (int j = 0; j < nblk; ++j) {
      const int st = j % kUpStages, use = j / kUpStages;
      if (use > 0) mbar_wait(&sm.kv_empty[st], (use - 1) & 1);
      const int t0 = j * kUpTiles;
      const int nt = min(kUpTiles, ntiles - t0);
      if (lane == 0) mbar_arrive_expect_tx(&sm.kv_full[st], static_cast<uint32_t>(nt * 2 * kTileBytesBf16));
      __syncwarp();
      // the block's 4 pages lie in one aligned group of 4 tiles, hence in one 32-page window of block ids
      const int first_page = t0 >> p.tpp_shift;
      if (first_page >= page0 + 32) {
        page0 = first_page & ~31;
        const int pg = page0 + lane;
        ids = (pg << p.tpp_shift) < ntiles ? __ldg(p.block_table + static_cast<int64_t>(b) * p.bt_stride + pg) : 0;
      }
      // 16 lanes issue the block's 16 boxes at once: lane = page * 4 + (V ? 2 : 0) + d half
      const int ti = (lane >> 2) & 3, half = lane & 1;
      const int tile = t0 + min(ti, nt - 1);
      const int blk = __shfl_sync(0xffffffffu, ids, (tile >> p.tpp_shift) - page0);
      if (lane < 16 && ti < nt) {
        const uint32_t bar = smem_u32(&sm.kv_full[st]);
        const int slot0 = (tile & p.tpp_mask) * kTileTokens;
        const uint32_t off = static_cast<uint32_t>(ti) * kHalfBytes;  // 16 rows of 128 B
        if (lane & 2)
          tma_load_5d_half(smem_u32(&sm.v[st][half][0]) + off, &tm_v, bar, 0, slot0, half, kvh, blk, pol);
        else
          tma_load_5d_half(smem_u32(&sm.k[st][half][0]) + off, &tm_k, bar, 0, slot0, half, kvh, blk, pol);
      }
    }
  } else if (warp == kMmaWarp) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, false, false);     // S = Q.K^T, B K-major
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 128, false, true);  // O += P.V, B MN-major
      auto issue_qk = [&](int j) {
        const int st = j % kUpStages;
        mbar_wait(&sm.kv_full[st], (j / kUpStages) & 1);
        tc_fence_after_sync();
        const uint32_t k_addr = smem_u32(&sm.k[st][0][0]);
#pragma unroll
        for (int t = 0; t < kUpQT; ++t) {
          const uint32_t q_addr = smem_u32(&sm.q[t][0][0]);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)  // d = 16 kk ..: half kk / 4, 32-byte step inside the swizzled row
            umma_bf16_ss(tmem + t * 128 + (j & 1) * 64,
                         umma_desc_sw128(q_addr + (kk >> 2) * kUpQHalf + (kk & 3) * 32, 16, 1024),
                         umma_desc_sw128(k_addr + (kk >> 2) * kUpKvHalf + (kk & 3) * 32, 16, 1024), idesc, kk > 0);
          umma_commit(&sm.s_full[t][j & 1]);
        }
      };
      if (nblk > 0) issue_qk(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) issue_qk(j + 1);
        const uint32_t v_addr = smem_u32(&sm.v[j % kUpStages][0][0]);
#pragma unroll
        for (int t = 0; t < kUpQT; ++t) {
          mbar_wait(&sm.p_full[t], j & 1);
          tc_fence_after_sync();
#if B200_UMMA_P_IN_TMEM
          const uint32_t p_tmem = tmem + t * 128 + (j & 1) * 64;  // P_t overwrote the first 32 columns of its S buffer
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)  // KV tokens 16 kk .. 16 kk + 15 = 8 TMEM columns of packed bf16 pairs
            umma_bf16_ts(tmem + 256 + t * 128, p_tmem + kk * 8, umma_desc_sw128(v_addr + kk * 2048, kUpKvHalf, 1024),
                         idesc_pv, j > 0 || kk > 0);
#else
          const uint32_t p_addr = smem_u32(&sm.p[t][0]);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)  // KV tokens 16 kk .. 16 kk + 15; N = 128: the two 64-wide MN atoms (d halves)
            umma_bf16_ss(tmem + 256 + t * 128, umma_desc_sw128(p_addr + kk * 32, 16, 1024),  // sit LBO = 8 KiB apart
                         umma_desc_sw128(v_addr + kk * 2048, kUpKvHalf, 1024), idesc_pv, j > 0 || kk > 0);
#endif
          umma_commit(&sm.pv_done[t]);
        }
        umma_commit(&sm.kv_empty[j % kUpStages]);  // every MMA that read this stage has been issued before
      }
    }
  } else {
    // ===================== softmax / correction / epilogue: warpgroup t, thread r = row r =====================
    const int t = warp >> 2, r = threadIdx.x & 127, i = t * kUpTok + (r >> 3), hh = r & 7;
    const bool row_ok = (i0 + i < q_len) && (hh < heads_here);
    const int pos = ctx0 + i0 + i;  // cache position of this row's query token
    const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tm_o = tmem + 256 + t * 128 + lane_off;
    // staged fp8 caches (b200attn_paged_prefill_fp8_staged): K and V hold the raw e4m3 values, the head's
    // dequantisation scales are folded into the softmax scale and the final normalisation
    const float scale = p.scale_log2 * (p.k_scale != nullptr ? __ldg(p.k_scale + kvh * p.scale_stride) : 1.f);
    float m = -CUDART_INF_F, l = 0.f;
    for (int j = 0; j < nblk; ++j) {
      const int buf = j & 1, st = j % kUpStages;
      const uint32_t tm_s = tmem + t * 128 + buf * 64 + lane_off;
      const int t_base = j * kUpBlk;
      const int lim = row_ok ? min(kv_end, pos + 1) - t_base : 0;  // columns [0, lim) of this block are visible
      const bool full = __all_sync(0xffffffffu, lim >= kUpBlk);    // no mask anywhere in this warp's rows
      mbar_wait(&sm.s_full[t][buf], (j >> 1) & 1);
      tc_fence_after_sync();
      // ---- the row's 64 scores: two TMEM loads in flight, one wait
      uint32_t x[2][32];
#pragma unroll
      for (int c = 0; c < 2; ++c) tmem_ld32(tm_s + c * 32, x[c]);
      tmem_ld_wait();
      float mx = -CUDART_INF_F;
      if (full) {
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int e = 0; e < 32; ++e) mx = fmaxf(mx, __uint_as_float(x[c][e]));
      } else {
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int e = 0; e < 32; ++e)
            if (c * 32 + e < lim) mx = fmaxf(mx, __uint_as_float(x[c][e]));
      }
      mx *= scale;  // scale > 0: max commutes with it
      const bool grow = mx > m + kTau;  // also the first visible block (m = -inf)
      const float m_new = grow ? mx : m;
      const float alpha = (grow && m != -CUDART_INF_F) ? fast_exp2(m - m_new) : 1.f;
      m = m_new;
      const float mu = (m == -CUDART_INF_F) ? 0.f : m;
      // ---- the P buffer is read by P.V of block j - 1 (issued the moment that P was published, long done by now);
      // the same completion also licenses the O rescale below
      if (j >= 1) {
        mbar_wait(&sm.pv_done[t], (j - 1) & 1);
        tc_fence_after_sync();
      }
      // ---- P = exp2(S * scale - m) as bf16 into the swizzled A-operand layout (one 128-byte row); row sum
      float sum = 0.f;
#if B200_UMMA_P_IN_TMEM
      uint32_t pk2[32];
#else
      uint8_t* prow = &sm.p[t][r * 128];
#endif
#pragma unroll
      for (int c = 0; c < 2; ++c) {
#if B200_UMMA_P_IN_TMEM
        uint32_t* pk = pk2 + 16 * c;
#else
        uint32_t pk[16];
#endif
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          float p0 = fast_exp2(fmaf(__uint_as_float(x[c][e]), scale, -mu));
          float p1 = fast_exp2(fmaf(__uint_as_float(x[c][e + 1]), scale, -mu));
          if (!full) {
            if (c * 32 + e >= lim) p0 = 0.f;
            if (c * 32 + e + 1 >= lim) p1 = 0.f;
          }
          sum += p0 + p1;
          pk[e >> 1] = pack_bf16x2(p0, p1);
        }
#if !B200_UMMA_P_IN_TMEM
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4)
          *reinterpret_cast<uint4*>(prow + (((c * 4 + q4) ^ (r & 7)) << 4)) =
              make_uint4(pk[4 * q4], pk[4 * q4 + 1], pk[4 * q4 + 2], pk[4 * q4 + 3]);
#endif
      }
#if B200_UMMA_P_IN_TMEM
      tmem_st32(tm_s, pk2);  // A operand of P.V straight from TMEM: row r = lane r, 32 columns of bf16 pairs
      tmem_st_wait();
#endif
      l = l * alpha + sum;
      // ---- V rows past the end of the chunk's context hold whatever the page (or an earlier block) held: zero them
      // so that 0 x NaN cannot reach O. Both warpgroups do it (same zeros) because each publishes only its own P.
      if (t_base + kUpBlk > kv_end && r < kUpBlk && t_base + r >= kv_end) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint4* vr = reinterpret_cast<uint4*>(&sm.v[st][half][r * 128]);
#pragma unroll
          for (int c = 0; c < 8; ++c) vr[c] = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      // ---- lazy rescale of O in TMEM (warp-uniform: tcgen05.ld/st are warp-wide); needs P.V of block j - 1 done
      if (j > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
#pragma unroll
        for (int c = 0; c < 4; c += 2) {
          tmem_ld32(tm_o + c * 32, x[0]);
          tmem_ld32(tm_o + c * 32 + 32, x[1]);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            x[0][e] = __float_as_uint(__uint_as_float(x[0][e]) * alpha);
            x[1][e] = __float_as_uint(__uint_as_float(x[1][e]) * alpha);
          }
          tmem_st32(tm_o + c * 32, x[0]);
          tmem_st32(tm_o + c * 32 + 32, x[1]);
        }
        tmem_st_wait();
      }
      fence_proxy_async_smem();
      tc_fence_before_sync();
      mbar_arrive(&sm.p_full[t]);
    }
    // ---- epilogue: O / l -> bf16 -> global (256 contiguous bytes per row)
    uint32_t x[32];
    const float inv = (l > 0.f ? 1.f / l : 0.f) * (p.v_scale != nullptr ? __ldg(p.v_scale + kvh * p.scale_stride) : 1.f);
    if (nblk > 0) {
      mbar_wait(&sm.pv_done[t], (nblk - 1) & 1);
      tc_fence_after_sync();
    }
    const int h = kvh * p.group + sub * kUpHeads + hh;
    uint4* dst = reinterpret_cast<uint4*>(p.out + static_cast<int64_t>(q0 + i0 + i) * p.out_stride_tok +
                                          static_cast<int64_t>(h) * kHeadDim);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if (nblk > 0) {
        tmem_ld32(tm_o + c * 32, x);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int e = 0; e < 32; ++e) x[e] = 0u;
      }
      if (row_ok) {
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          uint4 o;
          o.x = pack_bf16x2(__uint_as_float(x[8 * q4]) * inv, __uint_as_float(x[8 * q4 + 1]) * inv);
          o.y = pack_bf16x2(__uint_as_float(x[8 * q4 + 2]) * inv, __uint_as_float(x[8 * q4 + 3]) * inv);
          o.z = pack_bf16x2(__uint_as_float(x[8 * q4 + 4]) * inv, __uint_as_float(x[8 * q4 + 5]) * inv);
          o.w = pack_bf16x2(__uint_as_float(x[8 * q4 + 6]) * inv, __uint_as_float(x[8 * q4 + 7]) * inv);
          dst[c * 4 + q4] = o;
        }
      }
    }
  }
  // ---- teardown: every TMEM access has completed before the allocating warp frees it
  tc_fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp) tmem_free<kUpTmemCols>(tmem);
}

size_t prefill_umma_smem_bytes() { return sizeof(UpSmem) + 1024; }

cudaError_t launch_paged_prefill_umma(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const PrefillParams& p,
                                      cudaStream_t stream) {
  static thread_local int configured_dev = -1;
  int dev = 0;
  if (cudaError_t e = cudaGetDevice(&dev)) return e;
  const size_t smem = prefill_umma_smem_bytes();
  if (dev != configured_dev) {
    if (cudaError_t e = cudaFuncSetAttribute(paged_prefill_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem)))
      return e;
    configured_dev = dev;
  }
  const dim3 grid(static_cast<unsigned>(p.num_tokens / kUpCtaTok + p.batch),
                  static_cast<unsigned>(p.num_kv_heads * p.subgroups));
  paged_prefill_umma_kernel<<<grid, kUpThreads, smem, stream>>>(tm_k, tm_v, p);
  return cudaGetLastError();
}

}  // namespace b200
