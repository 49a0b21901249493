// Shared pieces of the split-K paged decode kernels (TMA variant and cp.async variant):
// parameter block, work-item decomposition, the per-stage tensor-core math on a swizzled
// shared-memory KV tile, and the cross-warp merge that ends a work item.
//
// Math orientation (GQA packing): the 8 query heads that share one KV head are the N
// dimension of mma.sync.m16n8k16 and the 16 KV tokens of a tile are the M dimension, i.e.
//   S^T[16 tok x 8 heads] = K[16 tok x 128] . Q^T[128 x 8 heads]
//   O^T[128 x 8 heads]   += V^T[128 x 16 tok] . P^T[16 tok x 8 heads]
// so no MMA row is padding (a [heads x tokens] orientation wastes half of M=16 at G=8).
// P^T is moved from the C-fragment layout to the B-fragment layout with movmatrix.trans.
#pragma once
#include <math_constants.h>

#include "ptx.cuh"

namespace b200 {

constexpr int kHeadDim = 128;      // only head_size 128 is built (all Qwen-Omni AR stages)
constexpr int kTileTokens = 16;    // one MMA M-tile of KV tokens
constexpr int kHeadsPerItem = 8;   // query heads packed in one N=8 tile
constexpr int kSkMaxBatch = 2048;  // stream-K plan (tiles prefix) lives in shared memory
constexpr int kSkWarps = 12;       // stream-K: warps per CTA, each with its own ring
constexpr int kSkDepth = 2;        // stream-K: ring stages per warp (8 KiB each)
constexpr int kSkDynPerWarp = 4;   // stream-K: dynamic (claimed) chunks per warp on average, at most
constexpr int kSkClaimLead = 2;    // stream-K: tiles left in the front's chunk when the next one is claimed
constexpr int kSkDynShiftDefault = 0;  // stream-K: dynamic share of the stream = 1 / 2^shift (0 = all static)
constexpr int kSk2PlanHdrInts = 16;     // stream-K v2 plan header (int32 words), see paged_decode_sk2.cu
constexpr int kSk2PlanMagic = 0x534b3250;  // "SK2P"
constexpr int kSk2FlagWaitFirst = 1;    // griddepcontrol.wait before the plan is read (plan made by the previous kernel)
constexpr int kSk2FlagEarlyKvRead = 2;  // KV tiles other than a sequence's last may be loaded before the wait
constexpr int kSk2FlagParity = 8;        // launch parity: consecutive launches on a stream alternate their claim words
constexpr int kSk3FlagEarlyQRead = 32;   // q may be loaded before griddepcontrol.wait (B200ATTN_DECODE_EARLY_Q_READ)
constexpr int kSk2MaxGroups = 512;      // CTAs of the persistent grid (= SMs) the control area has room for
constexpr int kSk2CalibMinTiles = 96;   // an SM's rate is only trusted if it streamed at least this many tiles
constexpr float kSk2CalibGain = 0.7f;   // share of the measured rate in a weight update
constexpr int kSk2CalibUpdatesDefault = 8;  // weight updates before the weights are frozen (B200ATTN_SK2_CALIB)
constexpr int kSk2FlagDebugInterleave = 4;  // debug: chunk c on CTA c % grid (timeline experiments; no merge, wrong output)
constexpr int kSk3CalibMinTiles = 3000;     // cooperative stream-K: an SM's rate is a sum over launches; it is used once it covers this many tiles
constexpr float kSk3CalibGainLate = 0.4f;   // cooperative stream-K: weight-update gain after the first four updates
constexpr int kSk3CalibUpdatesDefault = 24; // ... and the number of updates before the weights are frozen (B200ATTN_SK3_CALIB)
constexpr int kSk3Warps = 12;            // cooperative stream-K (paged_decode_sk3.cu): warps per CTA
constexpr int kSk3PlanHdrInts = 16;
constexpr int kSk3PlanMagic = 0x534b3350;  // "SK3P"
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// Control words of stream-K v2 at byte offset 256 of a decode workspace; all zero after
// b200attn_workspace_init, counters left at zero by every launch.
struct Sk2Ctrl {
  int32_t claimed[2][kSk2MaxGroups];  // chunk group claimed in the launch of parity 0 / 1
  int32_t finished[2];                // CTAs of the launch that are done
  int32_t calib_updates;              // weight updates made so far
  int32_t pad;
  float weight[kSk2MaxGroups];        // share of the tile stream per SM (0 everywhere: uniform)
  int32_t stat_tiles[kSk2MaxGroups];  // last launch: tiles the CTA on this SM streamed ...
  int32_t stat_cycles[kSk2MaxGroups]; // ... and the cycles it took
  int32_t claim3[kSk2MaxGroups];      // cooperative stream-K: id of the last launch of the step that took the group
};

struct DecodeParams {
  const __nv_bfloat16* q;
  int64_t q_stride_tok;
  const uint8_t* k_cache;  // bf16 or e4m3 bytes
  const uint8_t* v_cache;
  int64_t cs_block, cs_slot, cs_head;  // in cache elements
  __nv_bfloat16* out;
  int64_t out_stride_tok;
  float* lse;  // [batch, Hq] natural log, may be null
  const int32_t* block_table;
  int64_t bt_stride;
  const int32_t* seq_lens;
  int32_t batch, num_q_heads, num_kv_heads;
  int32_t group;         // Hq / Hkv
  int32_t subgroups;     // ceil(group / 8): head groups per KV head
  int32_t block_size;
  int32_t tiles_per_page;  // block_size / 16
  int32_t tpp_shift;       // log2(tiles_per_page) when it is a power of two, else -1
  int32_t tpp_mask;        // tiles_per_page - 1 (valid with tpp_shift >= 0)
  int32_t num_splits;
  int32_t tiles_per_split;  // multiple of the stage tile count
  int32_t total_items;      // batch * Hkv * num_splits * subgroups
  float scale_log2;         // softmax_scale * log2(e)
  const float* k_scale;
  const float* v_scale;
  int32_t scale_stride;  // 0 = per-tensor, 1 = per-kv-head
  float* part_o;         // [batch, Hq, num_splits, 128] f32, normalised partial outputs
  float* part_lse;       // [batch, Hq, num_splits] f32, natural log
  int32_t* plan_cum;     // stream-K only: [batch + 1] exclusive prefix of tiles per sequence
  int32_t* sk_ctrl;      // stream-K only: {claim counter, finished-warp counter}, zero between launches
  int32_t sk_dyn_shift;  // stream-K only: dynamic share of the stream = 1 / 2^shift (0 = all static)
  unsigned long long* trace;  // debug: per-warp {entry, static end, end} globaltimer stamps, or null
  // stream-K v2 (paged_decode_sk2.cu)
  const int32_t* sk2_plan;  // layout in paged_decode_sk2.cu
  Sk2Ctrl* sk2_ctrl;
  int32_t flags;            // kSk2Flag*
  // cooperative stream-K (paged_decode_sk3.cu): plan, partial slots [2 parities][2 * teams][8][128] (+ LSEs),
  // arrival counters [2 parities][teams]
  const int32_t* sk3_plan;
  float* sk3_part_o;
  float* sk3_part_lse;
  int32_t* sk3_cnt;
  int32_t launch_id;  // > 0, grows from one decode launch of a step to the next (claims, parity of the counter / slot sets)
};

// Stream-K schedule, computed identically by the decode and the reduce kernel from the tile total:
// the first S tiles are cut into nw static chunks of ts tiles (chunk id = global warp id); the
// remaining total - S tiles (about tpw >> dyn_shift per warp) are cut into nd dynamic chunks of cd
// tiles (chunk id = nw + j) that the warps claim from a global counter as they finish — SMs do not
// stream at equal rates (sm__cycles_active min/avg/max = 111k/140k/151k in the static-only capture),
// so the tail is balanced at run time.
struct SkSchedule {
  int64_t total;  // tiles x head-groups in the step
  int64_t S;      // static region [0, S)
  int ts;         // static tiles per warp
  int cd;         // tiles per dynamic chunk
  int nd;         // dynamic chunks
  bool dyn;
};
__host__ __device__ __forceinline__ SkSchedule sk_schedule(int64_t total, int nw, int dyn_shift) {
  SkSchedule sc;
  sc.total = total;
  const int64_t tpw64 = (total + nw - 1) / nw;
  const int tpw = static_cast<int>(tpw64 < 1 ? 1 : tpw64);
  int dpw = dyn_shift >= 1 ? (tpw >> dyn_shift) : 0;  // dynamic tiles per warp (average)
  if (dpw < 2) dpw = 0;
  sc.ts = tpw - dpw;
  const int64_t s_all = static_cast<int64_t>(nw) * sc.ts;
  sc.S = s_all < total ? s_all : total;
  int cd = (dpw + kSkDynPerWarp - 1) / kSkDynPerWarp;
  sc.cd = cd < 2 ? 2 : cd;
  sc.nd = static_cast<int>((total - sc.S + sc.cd - 1) / sc.cd);
  sc.dyn = dpw > 0 && sc.nd > 0;
  return sc;
}

struct WorkItem {
  int b, kvh, split, sub;
  int seq_len;
  int tile_begin, tile_end;  // 16-token tiles of this item
  int actual_splits;         // splits that own >= 1 tile for this sequence (>= 1)
};

// item index -> (b, kvh, split, sub); sub (head sub-group) is fastest so CTAs that re-read the
// same pages run next to each other in time.
__device__ __forceinline__ void item_coords(const DecodeParams& p, int item, WorkItem& w) {
  w.sub = item % p.subgroups;
  int r = item / p.subgroups;
  w.split = r % p.num_splits;
  r /= p.num_splits;
  w.kvh = r % p.num_kv_heads;
  w.b = r / p.num_kv_heads;
}
// Tile range of the item once the sequence length is known.
__device__ __forceinline__ void item_finish(const DecodeParams& p, int seq_len, WorkItem& w) {
  w.seq_len = seq_len < 0 ? 0 : seq_len;
  const int tiles_total = (w.seq_len + kTileTokens - 1) / kTileTokens;
  w.tile_begin = w.split * p.tiles_per_split;
  // the last split extends to the end of the sequence, so a sequence longer than the host's
  // max_seq_len hint is still fully covered
  w.tile_end = (w.split == p.num_splits - 1) ? tiles_total
                                             : min(tiles_total, w.tile_begin + p.tiles_per_split);
  if (w.tile_begin > w.tile_end) w.tile_begin = w.tile_end;
  int as = (tiles_total + p.tiles_per_split - 1) / p.tiles_per_split;
  w.actual_splits = max(1, min(as, p.num_splits));
}
__device__ __forceinline__ WorkItem decode_item(const DecodeParams& p, int item) {
  WorkItem w;
  item_coords(p, item, w);
  item_finish(p, __ldg(p.seq_lens + w.b), w);
  return w;
}

// Does this item produce output? Empty splits of short sequences produce nothing, except split 0
// of an empty sequence, which writes the zero row.
__device__ __forceinline__ bool item_active(const WorkItem& w) {
  return (w.tile_end > w.tile_begin) || (w.seq_len == 0 && w.split == 0);
}

// ------------------------------------------------------------------------------------------
// Shared-memory tile layout (identical to what TMA SWIZZLE_128B writes for a box of
// 64 elements x 16 rows): a 16-token x 128-dim bf16 tile is two 2 KiB halves (d 0..63, d 64..127);
// within a half row r (128 B) holds its eight 16-byte chunks at chunk index (c ^ (r & 7)).
// ------------------------------------------------------------------------------------------
constexpr int kHalfBytes = 16 * 128;            // 2048
constexpr int kTileBytesBf16 = 2 * kHalfBytes;  // 4096: one K (or V) tile
constexpr int kKvTileBytesBf16 = 2 * kTileBytesBf16;  // K tile followed by V tile

__device__ __forceinline__ uint32_t swz_off(int row, int chunk) {
  return static_cast<uint32_t>(row * 128 + ((chunk ^ (row & 7)) << 4));
}

// Per-lane ldmatrix byte offsets inside a half for the four (k-step mod 4) positions.
struct LaneOffsets {
  uint32_t k[4];  // K as A operand (row-major tokens x d)
  uint32_t v[4];  // V^T as A operand via ldmatrix.trans
};
__device__ __forceinline__ LaneOffsets make_lane_offsets(int lane) {
  LaneOffsets o;
  const int ri = lane & 7;
  {  // K: matrices (tok 0-7,c0) (tok 8-15,c0) (tok 0-7,c0+1) (tok 8-15,c0+1)
    const int row = ri + ((lane >> 3) & 1) * 8;
    const int cbit = lane >> 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) o.k[i] = swz_off(row, (i << 1) | cbit);
  }
  {  // V: matrices (tok 0-7,c0) (tok 0-7,c0+1) (tok 8-15,c0) (tok 8-15,c0+1), transposed on load
    const int row = ri + (lane >> 4) * 8;
    const int cbit = (lane >> 3) & 1;
#pragma unroll
    for (int i = 0; i < 4; ++i) o.v[i] = swz_off(row, (i << 1) | cbit);
  }
  return o;
}

// Running softmax state of one warp for one work item. Heads (2t, 2t+1) with t = lane & 3;
// m is in the log2 domain (score * scale * log2e); l is the lane-partial row sum.
struct WarpState {
  float m[2];
  float l[2];
  float acc[8][4];  // O^T: d-tile j, C-fragment (d=16j+g | +8, head 2t | 2t+1)
};
__device__ __forceinline__ void state_reset(WarpState& st) {
  st.m[0] = st.m[1] = -CUDART_INF_F;
  st.l[0] = st.l[1] = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j)
#pragma unroll
    for (int i = 0; i < 4; ++i) st.acc[j][i] = 0.f;
}

// Q^T B-fragments for the 8 k-steps over d: qf[kk][0] = Q[head g][16kk+2t, +1],
// qf[kk][1] = Q[head g][16kk+2t+8, +9]; heads beyond the group are zero.
__device__ __forceinline__ void load_q_frags(const DecodeParams& p, const WorkItem& w, int lane,
                                             uint32_t (&qf)[8][2]) {
  const int g = lane >> 2, t = lane & 3;
  const int head_in_group = w.sub * kHeadsPerItem + g;
  const bool valid = head_in_group < p.group;
  const int h = w.kvh * p.group + head_in_group;
  const uint32_t* row = reinterpret_cast<const uint32_t*>(
      p.q + static_cast<int64_t>(w.b) * p.q_stride_tok + static_cast<int64_t>(h) * kHeadDim);
#pragma unroll
  for (int kk = 0; kk < 8; ++kk) {
    qf[kk][0] = valid ? __ldg(row + kk * 8 + t) : 0u;
    qf[kk][1] = valid ? __ldg(row + kk * 8 + t + 4) : 0u;
  }
}

// Zero the V rows [valid, 16*NT) of a stage so that P=0 never multiplies non-finite garbage
// from the unused tail of a sequence's last page.
template <int NT>
__device__ __forceinline__ void zero_v_tail(uint8_t* stage, int valid, int lane) {
  const int chunk = lane & 7;
  for (int idx = (lane >> 3); idx < NT * 16 * 2; idx += 4) {  // (tile,row,half) triples
    const int ti = idx / 32, row = (idx >> 1) & 15, half = idx & 1;
    if (ti * 16 + row >= valid) {
      uint8_t* p = stage + ti * kKvTileBytesBf16 + kTileBytesBf16 + half * kHalfBytes + row * 128 +
                   chunk * 16;
      *reinterpret_cast<int4*>(p) = make_int4(0, 0, 0, 0);
    }
  }
  __syncwarp();
}

// One pipeline stage of NT (1 or 2) 16-token tiles: S^T = K Q^T, online softmax, O^T += V^T P^T.
// `stage` points at [tile][K|V][half][16 rows][128 B]; `valid` = number of valid tokens in the
// stage (tokens >= valid are masked).
template <int NT>
__device__ __forceinline__ void consume_stage_bf16(uint8_t* stage, int valid, bool need_mask,
                                                   const uint32_t (&qf)[8][2],
                                                   const LaneOffsets& lo, float scale_log2,
                                                   WarpState& st, int lane) {
  const uint32_t base = smem_u32(stage);
  float s[NT][4];
#pragma unroll
  for (int ti = 0; ti < NT; ++ti) s[ti][0] = s[ti][1] = s[ti][2] = s[ti][3] = 0.f;

#pragma unroll
  for (int kk = 0; kk < 8; ++kk) {
#pragma unroll
    for (int ti = 0; ti < NT; ++ti) {
      uint32_t a0, a1, a2, a3;
      ldsm_x4(base + ti * kKvTileBytesBf16 + (kk >> 2) * kHalfBytes + lo.k[kk & 3], a0, a1, a2, a3);
      mma_bf16_16816(s[ti], a0, a1, a2, a3, qf[kk][0], qf[kk][1]);
    }
  }

  const int g = lane >> 2;
  if (need_mask) {
#pragma unroll
    for (int ti = 0; ti < NT; ++ti) {
      if (ti * 16 + g >= valid) s[ti][0] = s[ti][1] = -CUDART_INF_F;
      if (ti * 16 + g + 8 >= valid) s[ti][2] = s[ti][3] = -CUDART_INF_F;
    }
    zero_v_tail<NT>(stage, valid, lane);
  }

  // row max per head over this stage's tokens
  float mx0 = fmaxf(s[0][0], s[0][2]), mx1 = fmaxf(s[0][1], s[0][3]);
#pragma unroll
  for (int ti = 1; ti < NT; ++ti) {
    mx0 = fmaxf(mx0, fmaxf(s[ti][0], s[ti][2]));
    mx1 = fmaxf(mx1, fmaxf(s[ti][1], s[ti][3]));
  }
#pragma unroll
  for (int off = 4; off <= 16; off <<= 1) {
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, off));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, off));
  }
  const float mn0 = fmaxf(st.m[0], mx0 * scale_log2);
  const float mn1 = fmaxf(st.m[1], mx1 * scale_log2);
  const float mu0 = (mn0 == -CUDART_INF_F) ? 0.f : mn0;  // all-masked / -inf rows stay NaN-free
  const float mu1 = (mn1 == -CUDART_INF_F) ? 0.f : mn1;
  const float al0 = fast_exp2(st.m[0] - mu0);
  const float al1 = fast_exp2(st.m[1] - mu1);
  st.m[0] = mn0;
  st.m[1] = mn1;

  uint32_t pb[NT][2];
  float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
  for (int ti = 0; ti < NT; ++ti) {
    const float p0 = fast_exp2(fmaf(s[ti][0], scale_log2, -mu0));
    const float p1 = fast_exp2(fmaf(s[ti][1], scale_log2, -mu1));
    const float p2 = fast_exp2(fmaf(s[ti][2], scale_log2, -mu0));
    const float p3 = fast_exp2(fmaf(s[ti][3], scale_log2, -mu1));
    sum0 += p0 + p2;
    sum1 += p1 + p3;
    pb[ti][0] = movmatrix_trans(pack_bf16x2(p0, p1));  // tokens 0-7  -> B-frag k 0-7
    pb[ti][1] = movmatrix_trans(pack_bf16x2(p2, p3));  // tokens 8-15 -> B-frag k 8-15
  }
  st.l[0] = fmaf(st.l[0], al0, sum0);
  st.l[1] = fmaf(st.l[1], al1, sum1);

  if (__any_sync(0xffffffffu, (al0 != 1.f) | (al1 != 1.f))) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      st.acc[j][0] *= al0;
      st.acc[j][1] *= al1;
      st.acc[j][2] *= al0;
      st.acc[j][3] *= al1;
    }
  }

#pragma unroll
  for (int j = 0; j < 8; ++j) {
#pragma unroll
    for (int ti = 0; ti < NT; ++ti) {
      uint32_t a0, a1, a2, a3;
      ldsm_x4_trans(base + ti * kKvTileBytesBf16 + kTileBytesBf16 + (j >> 2) * kHalfBytes +
                        lo.v[j & 3],
                    a0, a1, a2, a3);
      mma_bf16_16816(st.acc[j], a0, a1, a2, a3, pb[ti][0], pb[ti][1]);
    }
  }
}

// ------------------------------------------------------------------------------------------
// Cross-warp merge at the end of a work item.
// ------------------------------------------------------------------------------------------
constexpr int kMergeRow = kHeadDim + 4;  // padded row (floats): conflict-free C-fragment stores

template <int W>
struct MergeSmem {
  float o[W][kHeadsPerItem][kMergeRow];
  float m[W][kHeadsPerItem];
  float l[W][kHeadsPerItem];
};

// Called by all W consumer warps (W*32 threads, named barrier `bar_id`).
template <int W>
__device__ __forceinline__ void merge_and_store(const DecodeParams& p, const WorkItem& w,
                                                WarpState& st, MergeSmem<W>& ms, int warp, int lane,
                                                int bar_id) {
  const int g = lane >> 2, t = lane & 3;
  // full row sums: reduce the lane partials over the 8 token rows (lanes sharing t)
#pragma unroll
  for (int off = 4; off <= 16; off <<= 1) {
    st.l[0] += __shfl_xor_sync(0xffffffffu, st.l[0], off);
    st.l[1] += __shfl_xor_sync(0xffffffffu, st.l[1], off);
  }
  named_bar_sync(bar_id, W * 32);  // previous merge's readers are done with `ms`
  if (g == 0) {
    ms.m[warp][2 * t] = st.m[0];
    ms.m[warp][2 * t + 1] = st.m[1];
    ms.l[warp][2 * t] = st.l[0];
    ms.l[warp][2 * t + 1] = st.l[1];
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    ms.o[warp][2 * t][16 * j + g] = st.acc[j][0];
    ms.o[warp][2 * t + 1][16 * j + g] = st.acc[j][1];
    ms.o[warp][2 * t][16 * j + g + 8] = st.acc[j][2];
    ms.o[warp][2 * t + 1][16 * j + g + 8] = st.acc[j][3];
  }
  named_bar_sync(bar_id, W * 32);

  const int tid = warp * 32 + lane;
  const bool direct = (w.actual_splits == 1);
  // each thread: one head, 4 consecutive d
  for (int idx = tid; idx < kHeadsPerItem * (kHeadDim / 4); idx += W * 32) {
    const int head = idx / (kHeadDim / 4);
    const int d0 = (idx % (kHeadDim / 4)) * 4;
    const int head_in_group = w.sub * kHeadsPerItem + head;
    if (head_in_group >= p.group) continue;
    const int h = w.kvh * p.group + head_in_group;
    float M = -CUDART_INF_F;
#pragma unroll
    for (int ww = 0; ww < W; ++ww) M = fmaxf(M, ms.m[ww][head]);
    float denom = 0.f, v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
    if (M != -CUDART_INF_F) {
#pragma unroll
      for (int ww = 0; ww < W; ++ww) {
        const float mw = ms.m[ww][head];
        const float wgt = (mw == -CUDART_INF_F) ? 0.f : fast_exp2(mw - M);
        denom = fmaf(ms.l[ww][head], wgt, denom);
        const float4 ov = *reinterpret_cast<const float4*>(&ms.o[ww][head][d0]);
        v0 = fmaf(ov.x, wgt, v0);
        v1 = fmaf(ov.y, wgt, v1);
        v2 = fmaf(ov.z, wgt, v2);
        v3 = fmaf(ov.w, wgt, v3);
      }
    }
    const float inv = (denom > 0.f) ? 1.f / denom : 0.f;
    v0 *= inv, v1 *= inv, v2 *= inv, v3 *= inv;
    const float lse_nat = (denom > 0.f) ? (M + log2f(denom)) * kLn2 : -CUDART_INF_F;
    if (direct) {
      __nv_bfloat16* dst = p.out + static_cast<int64_t>(w.b) * p.out_stride_tok +
                           static_cast<int64_t>(h) * kHeadDim + d0;
      uint2 pk;
      pk.x = pack_bf16x2(v0, v1);
      pk.y = pack_bf16x2(v2, v3);
      *reinterpret_cast<uint2*>(dst) = pk;
      if (p.lse != nullptr && d0 == 0) p.lse[static_cast<int64_t>(w.b) * p.num_q_heads + h] = lse_nat;
    } else {
      const int64_t row = (static_cast<int64_t>(w.b) * p.num_q_heads + h) * p.num_splits + w.split;
      *reinterpret_cast<float4*>(p.part_o + row * kHeadDim + d0) = make_float4(v0, v1, v2, v3);
      if (d0 == 0) p.part_lse[row] = lse_nat;
    }
  }
}

}  // namespace b200
