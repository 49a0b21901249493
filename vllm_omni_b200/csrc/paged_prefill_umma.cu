// Causal varlen attention over the paged KV cache on the 5th-generation tensor cores (tcgen05 / UMMA,
// accumulators in TMEM) — the "dense enough to pay off" case of GQA query-group packing: a CTA owns 16 query
// tokens x the 8 query heads of one KV head = 128 rows, exactly one M = 128 UMMA tile with no padding at G = 8
// (the mma.sync kernel in paged_prefill.cu stays the path for FP8 caches and is the comparison point).
// Same call as paged_prefill.cu (vllm/v1/attention/backends/flash_attn.py:758-819 with max_query_len > 1).
//
// Per KV block of 128 tokens (8 pages of 16):
//   S[128 x 128]  = Q[128 x 128d] . K^T      8 x tcgen05.mma m128 n128 k16, A and B K-major from shared memory
//   P = exp2(S * scale - m) (causal / tail masked), row statistics in registers: thread r owns row r (TMEM lane r)
//   O[128 x 128d] += P[128 x 128] . V        16 x tcgen05.mma m128 n64 k16, A = P (bf16, written to shared memory by
//                                            the softmax threads in the swizzled K-major layout), B = V MN-major
// K/V pages are staged by TMA straight into the SWIZZLE_128B layout the UMMA descriptors describe (two boxes of
// 64 d x 16 tokens per page, one per 128-byte half of the head dimension), two stages of 64 KiB. S is double-
// buffered in TMEM (2 x 128 columns) and P in shared memory, so Q.K^T of block j+1 and P.V of block j-1 run under
// the softmax of block j; O lives in 128 more columns and is rescaled in TMEM only when a row maximum grows by
// more than 2^kTau (lazy rescale).
//
// Warp roles (192 threads): warps 0-3 softmax / correction / epilogue (warp w touches TMEM lanes 32w..32w+31),
// warp 4 TMA producer, warp 5 TMEM allocation + single-thread MMA issue.
#include "kernels.h"
#include "sk_math.cuh"
#include "umma.cuh"

namespace b200 {

constexpr int kUpTok = 16;             // query tokens per CTA
constexpr int kUpHeads = 8;            // query-head slots per CTA (rows = token * 8 + head slot)
constexpr int kUpBlk = 128;            // KV tokens per block
constexpr int kUpTiles = kUpBlk / kTileTokens;  // 16-token pages per block
constexpr int kUpHalf = 128 * 128;     // one [128 rows][128 B] swizzled operand half: 16 KiB
constexpr int kUpStages = 2;
constexpr int kUpThreads = 192;
constexpr uint32_t kUpTmemCols = 512;  // S0 | S1 | O | (unused): power of two

struct UpSmem {
  alignas(1024) uint8_t q[2][kUpHalf];             // A of Q.K^T: [d half][row][64 d]
  alignas(1024) uint8_t p[2][2][kUpHalf];          // A of P.V, double-buffered: [buffer][token half][row][64 tokens]
  alignas(1024) uint8_t k[kUpStages][2][kUpHalf];  // B of Q.K^T: [d half][token][64 d]
  alignas(1024) uint8_t v[kUpStages][2][kUpHalf];  // B of P.V (MN-major): [d half][token][64 d]
  alignas(8) uint64_t kv_full[kUpStages];
  alignas(8) uint64_t kv_empty[kUpStages];
  alignas(8) uint64_t s_full[2];
  alignas(8) uint64_t p_full;
  alignas(8) uint64_t pv_done[2];  // P.V of block j has completed: barrier j & 1 (never more than one phase ahead of a waiter)
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

// same tile-slot scheme as paged_prefill.cu (pf_locate)
__device__ __forceinline__ bool up_locate(const PrefillParams& p, int slot, int& b, int& qt, int& q0, int& q_len) {
  int lo = 0, hi = p.batch;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if ((__ldg(p.query_start_loc + mid) >> 4) + mid <= slot) lo = mid; else hi = mid;
  }
  b = lo;
  q0 = __ldg(p.query_start_loc + b);
  q_len = __ldg(p.query_start_loc + b + 1) - q0;
  qt = slot - ((q0 >> 4) + b);
  return qt * kUpTok < q_len;
}

__global__ void __launch_bounds__(kUpThreads, 1)
paged_prefill_umma_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                          const PrefillParams p) {
  extern __shared__ uint8_t smem_raw[];
  UpSmem& sm = *reinterpret_cast<UpSmem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  int b, qt, q0, q_len;
  // heavy (late) tiles first; the whole CTA leaves together, before any barrier or TMEM allocation
  if (!up_locate(p, static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x), b, qt, q0, q_len)) return;
  const int hs = blockIdx.y;
  const int kvh = hs / p.subgroups, sub = hs % p.subgroups;
  const int heads_here = min(kUpHeads, p.group - sub * kUpHeads);
  const int S = max(0, __ldg(p.seq_lens + b));
  const int i0 = qt * kUpTok;
  const int ctx0 = S - q_len;  // cache position of the sequence's first new token
  const int kv_end = max(0, min(S, ctx0 + i0 + kUpTok));
  const int ntiles = (kv_end + kTileTokens - 1) / kTileTokens;
  const int nblk = (kv_end + kUpBlk - 1) / kUpBlk;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kUpStages; ++s) {
      mbar_init(&sm.kv_full[s], 1);
      mbar_init(&sm.kv_empty[s], 1);
      mbar_init(&sm.s_full[s], 1);
    }
    mbar_init(&sm.p_full, 128);
    mbar_init(&sm.pv_done[0], 1);
    mbar_init(&sm.pv_done[1], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
  }
  if (warp == 5) tmem_alloc<kUpTmemCols>(&sm.tmem_base);
  // ---- Q tile -> shared memory (swizzled K-major A operand); rows of absent tokens / head slots are zero
  if (warp < 4) {
    const int r = threadIdx.x, i = r >> 3, hh = r & 7;
    const bool ok = (i0 + i < q_len) && (hh < heads_here);
    const int h = kvh * p.group + sub * kUpHeads + hh;
    const int4* src = reinterpret_cast<const int4*>(p.q + static_cast<int64_t>(q0 + i0 + i) * p.q_stride_tok +
                                                    static_cast<int64_t>(h) * kHeadDim);
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      const int4 x = ok ? __ldg(src + c) : make_int4(0, 0, 0, 0);
      *reinterpret_cast<int4*>(&sm.q[c >> 3][r * 128 + (((c & 7) ^ (r & 7)) << 4)]) = x;
    }
    fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's async-proxy reads
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = sm.tmem_base;
  const uint32_t tm_s0 = tmem, tm_o = tmem + 256;

  if (warp == 4) {
    // ===================== TMA producer =====================
    const uint64_t pol = l2_policy_evict_last();  // the other query tiles of the sequence re-read these pages
    int page0 = -32, ids = 0;
    for (int j = 0; j < nblk; ++j) {
      const int st = j & 1;
      if (j >= kUpStages) mbar_wait(&sm.kv_empty[st], ((j >> 1) - 1) & 1);
      const int t0 = j * kUpTiles;
      const int nt = min(kUpTiles, ntiles - t0);
      if (lane == 0) mbar_arrive_expect_tx(&sm.kv_full[st], static_cast<uint32_t>(nt * 2 * kTileBytesBf16));
      for (int ti = 0; ti < nt; ++ti) {
        const int tile = t0 + ti;
        const int page = tile >> p.tpp_shift;
        if (page >= page0 + 32) {
          page0 = page & ~31;
          const int pg = page0 + lane;
          ids = (pg << p.tpp_shift) < ntiles ? __ldg(p.block_table + static_cast<int64_t>(b) * p.bt_stride + pg) : 0;
        }
        const int blk = __shfl_sync(0xffffffffu, ids, page - page0);
        if (lane == 0) {
          const uint32_t bar = smem_u32(&sm.kv_full[st]);
          const int slot0 = (tile & p.tpp_mask) * kTileTokens;
          const uint32_t off = static_cast<uint32_t>(ti) * kHalfBytes;  // 16 rows of 128 B
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            tma_load_5d_half(smem_u32(&sm.k[st][half][0]) + off, &tm_k, bar, 0, slot0, half, kvh, blk, pol);
            tma_load_5d_half(smem_u32(&sm.v[st][half][0]) + off, &tm_v, bar, 0, slot0, half, kvh, blk, pol);
          }
        }
      }
    }
  } else if (warp == 5) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      constexpr uint32_t idesc_qk = umma_idesc_bf16(128, 128, false, false);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, false, true);
      const uint32_t q_addr = smem_u32(&sm.q[0][0]);
      auto issue_qk = [&](int j) {
        const int st = j & 1;
        mbar_wait(&sm.kv_full[st], (j >> 1) & 1);
        tc_fence_after_sync();
        const uint32_t k_addr = smem_u32(&sm.k[st][0][0]);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {  // d = 16 kk .. 16 kk + 15: half kk / 4, 32-byte step inside the swizzled row
          const uint32_t o = (kk >> 2) * kUpHalf + (kk & 3) * 32;
          umma_bf16_ss(tm_s0 + st * 128, umma_desc_sw128(q_addr + o, 16, 1024), umma_desc_sw128(k_addr + o, 16, 1024),
                       idesc_qk, kk > 0);
        }
        umma_commit(&sm.s_full[st]);
      };
      if (nblk > 0) issue_qk(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) issue_qk(j + 1);
        mbar_wait(&sm.p_full, j & 1);
        tc_fence_after_sync();
        const uint32_t v_addr = smem_u32(&sm.v[j & 1][0][0]), p_addr = smem_u32(&sm.p[j & 1][0][0]);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {  // KV tokens 16 kk .. 16 kk + 15
          const uint64_t da = umma_desc_sw128(p_addr + (kk >> 2) * kUpHalf + (kk & 3) * 32, 16, 1024);
#pragma unroll
          for (int nh = 0; nh < 2; ++nh)  // d halves: one 64-wide MN atom each
            umma_bf16_ss(tm_o + nh * 64, da, umma_desc_sw128(v_addr + nh * kUpHalf + kk * 2048, kUpHalf, 1024),
                         idesc_pv, j > 0 || kk > 0);
        }
        umma_commit(&sm.kv_empty[j & 1]);
        umma_commit(&sm.pv_done[j & 1]);
      }
    }
  } else {
    // ===================== softmax / correction / epilogue: thread r = row r =====================
    const int r = threadIdx.x, i = r >> 3, hh = r & 7;
    const bool row_ok = (i0 + i < q_len) && (hh < heads_here);
    const int pos = ctx0 + i0 + i;  // cache position of this row's query token
    const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
    const float scale = p.scale_log2;
    float m = -CUDART_INF_F, l = 0.f;
    for (int j = 0; j < nblk; ++j) {
      const int st = j & 1;
      const uint32_t tm_s = tm_s0 + st * 128 + lane_off;
      const int t_base = j * kUpBlk;
      const int lim = row_ok ? min(kv_end, pos + 1) - t_base : 0;  // columns [0, lim) of this block are visible
      const bool full = __all_sync(0xffffffffu, lim >= kUpBlk);    // no mask anywhere in this warp's rows
      mbar_wait(&sm.s_full[st], (j >> 1) & 1);
      tc_fence_after_sync();
      // ---- the row's 128 scores: four TMEM loads in flight, one wait
      uint32_t x[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32(tm_s + c * 32, x[c]);
      tmem_ld_wait();
      float mx = -CUDART_INF_F;
      if (full) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int e = 0; e < 32; ++e) mx = fmaxf(mx, __uint_as_float(x[c][e]));
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c)
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
      // ---- this P buffer was last read by P.V of block j - 2
      if (j >= 2) mbar_wait(&sm.pv_done[st], ((j >> 1) - 1) & 1);
      // ---- P = exp2(S * scale - m) as bf16 into the swizzled A-operand layout; row sum
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t pk[16];
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
        uint8_t* prow = &sm.p[st][c >> 1][r * 128];
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4)
          *reinterpret_cast<uint4*>(prow + ((((c & 1) * 4 + q4) ^ (r & 7)) << 4)) =
              make_uint4(pk[4 * q4], pk[4 * q4 + 1], pk[4 * q4 + 2], pk[4 * q4 + 3]);
      }
      l = l * alpha + sum;
      // ---- V rows past the end of the sequence hold whatever the page (or an earlier block) held: zero them so
      // that 0 x NaN cannot reach O (thread r owns token row r of the block)
      if (t_base + kUpBlk > kv_end && t_base + r >= kv_end) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint4* vr = reinterpret_cast<uint4*>(&sm.v[st][half][r * 128]);
#pragma unroll
          for (int c = 0; c < 8; ++c) vr[c] = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      // ---- lazy rescale of O in TMEM (warp-uniform: tcgen05.ld/st are warp-wide); needs P.V of block j - 1 done
      if (j > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
        mbar_wait(&sm.pv_done[st ^ 1], ((j - 1) >> 1) & 1);
        tc_fence_after_sync();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          tmem_ld32(tm_o + lane_off + c * 32, x[c]);
        }
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
#pragma unroll
          for (int e = 0; e < 32; ++e) x[c][e] = __float_as_uint(__uint_as_float(x[c][e]) * alpha);
          tmem_st32(tm_o + lane_off + c * 32, x[c]);
        }
        tmem_st_wait();
      }
      fence_proxy_async_smem();
      tc_fence_before_sync();
      mbar_arrive(&sm.p_full);
    }
    uint32_t x[32];
    // ---- epilogue: O / l -> bf16 -> global (256 contiguous bytes per row)
    const float inv = l > 0.f ? 1.f / l : 0.f;
    if (nblk > 0) {
      mbar_wait(&sm.pv_done[(nblk - 1) & 1], ((nblk - 1) >> 1) & 1);
      tc_fence_after_sync();
    }
    const int h = kvh * p.group + sub * kUpHeads + hh;
    uint4* dst = reinterpret_cast<uint4*>(p.out + static_cast<int64_t>(q0 + i0 + i) * p.out_stride_tok +
                                          static_cast<int64_t>(h) * kHeadDim);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if (nblk > 0) {
        tmem_ld32(tm_o + lane_off + c * 32, x);
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
  if (warp == 5) tmem_free<kUpTmemCols>(tmem);
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
  const dim3 grid(static_cast<unsigned>(p.num_tokens / kUpTok + p.batch),
                  static_cast<unsigned>(p.num_kv_heads * p.subgroups));
  paged_prefill_umma_kernel<<<grid, kUpThreads, smem, stream>>>(tm_k, tm_v, p);
  return cudaGetLastError();
}

}  // namespace b200
