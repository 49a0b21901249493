// Paged decode attention for sm_100a — cooperative stream-K (third generation, the default): ONE kernel
// per layer, no reduce launch.
//
// What the per-warp timeline of the second generation showed (profiles/r2_baseline_sk2.md): with one
// contiguous chunk per WARP the 12 warps of an SM end 6 us apart on average (an SM streams with fewer and
// fewer warps at the end), every item is cut into ~8 partials that travel through global memory, and the
// log-sum-exp reduce launch that folds them costs 5 us per layer. This kernel changes the unit of the cut:
//
//  * The tile stream of the step — all 16-token KV tiles ordered (sequence, kv head [, 8-head sub-group],
//    tile) — is cut once per CTA (148 cuts, sized by the SM's measured rate like before), and a CTA range is
//    worked on by TEAMS of T warps (T divides 12; T = 12 for long contexts, smaller when items are short,
//    chosen by the plan kernel from the step's average item length). The warps of a team take the positions
//    of the team's range round-robin (position p -> warp p mod T), so they walk through the same item at the
//    same time, end within one tile of each other, and fold their T running states in SHARED memory at the
//    end of every item (two named barriers per item, no global traffic).
//  * Only an item cut by a team boundary (<= one per team) leaves a partial in global memory, and no kernel
//    is launched to fold it: the team that holds the START of the item (its last piece of work) is the
//    item's owner; the teams that hold the rest reach their piece FIRST, i.e. their partials are written at
//    the start of the kernel. When the owner gets to the item at the end of its run it finds the arrival
//    counter complete, loads the neighbours' 4 KiB partials next to its own shared-memory merge and writes
//    the final row. If a neighbour is late (an item longer than two ranges) the protocol degrades to
//    "last arriver merges" through one atomic per piece — nobody ever spins on another CTA.
//  * Everything else is the second generation's: per-warp private ring of TMA stages issued by the warp's
//    own lane 0, tensor-core tile math of sk_math.cuh, plan once per step, early KV reads under programmatic
//    dependent launch, chunk group claimed by SM id (with the launch's id: nothing to re-arm on the way out).
//
// Replaces the decode branch of the reference's stock backends
// (vllm/v1/attention/backends/flashinfer.py:1694-1822, flash_attn.py:758-819) and their split-KV
// combine (vllm_flash_attn flash_fwd_combine; vllm/v1/attention/ops/triton_merge_attn_states.py).
#include "kernels.h"
#include "sk_math.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------
// Plan layout (int32 words), G = CTAs of the grid, W = 12, NTmax = G * W teams at most:
//   header[16] | tstart[NTmax + 1] | tdesc[NTmax][4] | tinfo[NTmax][4] | wdesc[G * W][4] | wids[G * W][32] | cum[batch + 1]
//   header     {positions, G, T, NT = G * (W / T), batch, hs_count, tpp_shift, magic, W / T}
//   tstart[c]  first stream position of team c (team c = CTA group c / (W/T), warps (c % (W/T)) * T ...)
//   tdesc[c]   {sequence, head-group, tile, tiles of the sequence} at tstart[c]
//   tinfo[c]   {owner team of the item the team's FIRST piece continues (-1: the range starts an item),
//               pieces of that item, teams that continue the team's LAST item when the team owns it and
//               it is cut (0: not cut), 1 if those are exactly the next teams (no empty team between)}
//   wdesc[w]   tdesc of the first position of warp w of the grid (sequence = batch: the warp has none)
//   wids[w]    the block ids of warp w's first 32 tiles of its first item (tile + T * lane): its first TMA needs no
//              dependent load
//   cum[b]     exclusive prefix of tiles per sequence
// ------------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ int64_t sk3_plan_tstart_off() { return kSk3PlanHdrInts; }
__host__ __device__ __forceinline__ int64_t sk3_plan_tdesc_off(int ntmax) { return kSk3PlanHdrInts + ((ntmax + 1 + 3) & ~3); }
__host__ __device__ __forceinline__ int64_t sk3_plan_tinfo_off(int ntmax) { return sk3_plan_tdesc_off(ntmax) + 4ll * ntmax; }
__host__ __device__ __forceinline__ int64_t sk3_plan_wdesc_off(int ntmax) { return sk3_plan_tdesc_off(ntmax) + 8ll * ntmax; }
__host__ __device__ __forceinline__ int64_t sk3_plan_wids_off(int ntmax) { return sk3_plan_tdesc_off(ntmax) + 12ll * ntmax; }
__host__ __device__ __forceinline__ int64_t sk3_plan_cum_off(int ntmax) { return sk3_plan_tdesc_off(ntmax) + 44ll * ntmax; }

size_t sk3_plan_ints(int batch, int groups) {
  return static_cast<size_t>(sk3_plan_cum_off(groups * kSk3Warps)) + ((static_cast<size_t>(batch) + 1 + 3) & ~size_t(3));
}
int decode_sk3_warps() { return kSk3Warps; }

// team that holds stream position t: the last c in [0, nt) with tstart[c] <= t
__device__ __forceinline__ int sk3_team_of(const int32_t* tstart, int nt, int t) {
  int lo = 0, hi = nt;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (tstart[mid] <= t) lo = mid; else hi = mid;
  }
  return lo;
}

// stream position -> {sequence, head-group, tile, tiles of the sequence}; cum[b] <= pos / hs_count < cum[b + 1]
__device__ __forceinline__ int4 sk3_desc_of(const int32_t* cum, int batch, int hs_count, int pos) {
  const int key = pos / hs_count;
  int lo = 0, hi = batch;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (cum[mid] <= key) lo = mid; else hi = mid;
  }
  const int n = cum[lo + 1] - cum[lo];
  const int rem = pos - cum[lo] * hs_count;
  return make_int4(lo, rem / n, rem % n, n);
}

constexpr int kSk3PlanSmemBatch = 4096;  // cum[] of batches up to this size is searched in shared memory

__device__ __forceinline__ float sk3_block_sum(float v, float* s_red, int tid) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  __syncthreads();  // s_red may still be read from the previous call
  if ((tid & 31) == 0) s_red[tid >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int w = 0; w < 32; ++w) t += s_red[w];
  return t;
}

// Team size for a step whose items average `avg` tiles: the largest divisor of 12 that still leaves every warp
// of a team about 8 tiles of an item (a merge costs two barriers and a pass over T states).
__host__ __device__ __forceinline__ int sk3_team_size(int avg) {
  if (avg >= 96) return 12;
  if (avg >= 48) return 6;
  if (avg >= 32) return 4;
  if (avg >= 24) return 3;
  if (avg >= 16) return 2;
  return 1;
}

// One CTA, once per step. Phase 0: SM weights from the rates the decode kernel recorded (frozen after
// `calib_max` updates, then a launch is a deterministic function of its inputs). Phase 1: tiles prefix.
// Phase 2: CTA ranges by weight, team ranges inside, team descriptors. Phase 3: warp descriptors.
// Phase 4: which teams share a cut item.
__global__ void __launch_bounds__(1024)
sk3_plan_kernel(const int32_t* __restrict__ seq_lens, int batch, int hs_count, int groups,
                const int32_t* __restrict__ block_table, int64_t bt_stride, int tpp_shift, int32_t* __restrict__ plan,
                Sk2Ctrl* __restrict__ ctrl, int calib_max, int team_force) {
  __shared__ int s_warp[32];
  __shared__ float s_red[32];
  __shared__ int s_carry, s_nonempty;
  __shared__ float s_w[kSk2MaxGroups];
  __shared__ int s_gstart[kSk2MaxGroups + 1];
  __shared__ int s_cum[kSk3PlanSmemBatch + 1];
  __shared__ int s_tstart[kSk2MaxGroups * kSk3Warps + 1];
  constexpr int W = kSk3Warps;
  const int ntmax = groups * W;
  int32_t* tstart = plan + sk3_plan_tstart_off();
  int4* tdesc = reinterpret_cast<int4*>(plan + sk3_plan_tdesc_off(ntmax));
  int4* tinfo = reinterpret_cast<int4*>(plan + sk3_plan_tinfo_off(ntmax));
  int4* wdesc = reinterpret_cast<int4*>(plan + sk3_plan_wdesc_off(ntmax));
  int32_t* cum = plan + sk3_plan_cum_off(ntmax);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    s_carry = 0;
    s_nonempty = 0;
  }
  if (tid < kSk2MaxGroups) ctrl->claim3[tid] = 0;  // the step's decode launches claim with ids > 0, growing
  // ---- phase 0: weights (thread g <-> SM g)
  {
    const bool have = tid < groups;
    const int tiles = have ? ctrl->stat_tiles[tid] : 0, cyc = have ? ctrl->stat_cycles[tid] : 0;
    const float w_raw = have ? ctrl->weight[tid] : 0.f;
    // (a step whose plan is made per call, or a small one, keeps summing until the sample is worth using)
    const float rate = (tiles >= kSk3CalibMinTiles && cyc > 0) ? static_cast<float>(tiles) / static_cast<float>(cyc) : 0.f;
    const int missing = __syncthreads_or(have && rate == 0.f);  // some SM has no (or too little) history
    const bool upd = !missing && ctrl->calib_updates < calib_max;
    const float rsum = sk3_block_sum(rate, s_red, tid), wsum = sk3_block_sum(w_raw, s_red, tid);
    const float w_old = wsum > 0.f ? w_raw / wsum : 1.f / groups;
    // the first updates follow the measurement closely, the later ones average the launch-to-launch noise out
    const float gain = ctrl->calib_updates < 4 ? kSk2CalibGain : kSk3CalibGainLate;
    float w = upd ? (1.f - gain) * w_old + gain * rate / rsum : w_old;
    if (!have) w = 0.f;
    const float norm = sk3_block_sum(w, s_red, tid);
    w /= norm;
    if (have) {
      s_w[tid] = w;
      if (upd) {
        ctrl->weight[tid] = w;
        ctrl->stat_tiles[tid] = 0;  // a history is used once
        ctrl->stat_cycles[tid] = 0;
      }
    }
    __syncthreads();
    if (tid == 0 && upd) ctrl->calib_updates += 1;
  }
  // ---- phase 1
  for (int base = 0; base < batch; base += 1024) {
    const int i = base + tid;
    const int s = i < batch ? max(0, seq_lens[i]) : 0;
    const int t = (s + kTileTokens - 1) / kTileTokens;
    int incl = t;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += v;
    }
    const unsigned nz = __ballot_sync(0xffffffffu, t > 0);
    if (lane == 31) s_warp[warp] = incl;
    if (lane == 0 && nz) atomicAdd(&s_nonempty, __popc(nz));
    __syncthreads();
    int wbase = 0;
    for (int w = 0; w < warp; ++w) wbase += s_warp[w];
    const int carry = s_carry;
    if (i < batch) {
      cum[i] = carry + wbase + incl - t;
      if (i < kSk3PlanSmemBatch) s_cum[i] = carry + wbase + incl - t;
    }
    __syncthreads();
    if (tid == 1023) s_carry = carry + wbase + incl;
    __syncthreads();
  }
  const int tiles = s_carry;
  if (static_cast<int64_t>(tiles) * hs_count > 0x7fffffffll) __trap();  // stream positions are 32-bit
  const int total = tiles * hs_count;
  const bool cum_in_smem = batch <= kSk3PlanSmemBatch;
  const int32_t* cum_r = cum_in_smem ? s_cum : cum;
  const int avg = tiles / max(1, s_nonempty);
  const int T = (team_force >= 1 && team_force <= W && W % team_force == 0) ? team_force : sk3_team_size(avg);
  const int tpc = W / T, nt = groups * tpc;
  if (tid == 0) {
    cum[batch] = tiles;
    if (cum_in_smem) s_cum[batch] = tiles;
    plan[0] = total;
    plan[1] = groups;
    plan[2] = T;
    plan[3] = nt;
    plan[4] = batch;
    plan[5] = hs_count;
    plan[6] = 0;
    plan[7] = kSk3PlanMagic;
    plan[8] = tpc;
    plan[9] = 0;
    plan[10] = 0;
    plan[11] = ctrl->calib_updates < calib_max ? 1 : 0;  // the decode launches of this step record their SMs' rates
    // CTA ranges: cumulative weights; every CTA gets at least one position while there are any, so the non-empty
    // CTAs are a prefix of the grid
    double acc = 0.0;
    s_gstart[0] = 0;
    for (int g = 0; g < groups; ++g) {
      acc += s_w[g];
      int e = static_cast<int>(acc * total + 0.5);
      e = min(max(e, s_gstart[g] + 1), total);
      s_gstart[g + 1] = g + 1 == groups ? total : e;
    }
  }
  __syncthreads();  // cum[] and s_gstart[] are read below
  // ---- phase 2: team ranges (a CTA range cut into W / T parts, the first (len % tpc) one position longer)
  for (int c = tid; c <= nt; c += 1024) {
    const int g = c / tpc, k = c % tpc;
    const int gs = s_gstart[min(g, groups)], ge = s_gstart[min(g + 1, groups)];
    const int q = (ge - gs) / tpc, r = (ge - gs) % tpc;
    const int t0 = c == nt ? total : gs + k * q + min(k, r);
    tstart[c] = t0;
    s_tstart[c] = t0;
    if (c < nt) tdesc[c] = t0 < total ? sk3_desc_of(cum_r, batch, hs_count, t0) : make_int4(batch, 0, 0, 1);
  }
  __syncthreads();
  // ---- phase 3: first position of every warp
  for (int idx = tid; idx < groups * W; idx += 1024) {
    const int g = idx / W, w = idx % W;
    const int c = g * tpc + w / T;
    const int pos = s_tstart[c] + w % T;
    wdesc[idx] = pos < s_tstart[c + 1] ? sk3_desc_of(cum_r, batch, hs_count, pos) : make_int4(batch, 0, 0, 1);
  }
  __syncthreads();
  // ---- phase 3b: the block ids of every warp's first tiles
  int32_t* wids = plan + sk3_plan_wids_off(ntmax);
  for (int idx = tid; idx < groups * W * 32; idx += 1024) {
    const int4 d = wdesc[idx >> 5];
    const int tile = d.z + T * (idx & 31);
    wids[idx] = (d.x < batch && tile < d.w) ? block_table[static_cast<int64_t>(d.x) * bt_stride + (tile >> tpp_shift)] : 0;
  }
  // ---- phase 4: cut items
  for (int c = tid; c < nt; c += 1024) {
    const int t0 = s_tstart[c], t1 = s_tstart[c + 1];
    int4 info = make_int4(-1, 0, 0, 1);
    if (t1 > t0) {
      const int4 d = sk3_desc_of(cum_r, batch, hs_count, t0);
      if (d.z > 0) {  // the range starts inside an item
        const int i0 = t0 - d.z, i1 = i0 + d.w;
        const int own = sk3_team_of(s_tstart, nt, i0);
        int np = 1;
        for (int j = own + 1; j < nt && s_tstart[j] < i1; ++j) np += s_tstart[j + 1] > s_tstart[j] ? 1 : 0;
        info.x = own;
        info.y = np | (sk3_team_of(s_tstart, nt, i1 - 1) - own == np - 1 ? (1 << 30) : 0);
      }
      const int4 e = sk3_desc_of(cum_r, batch, hs_count, t1 - 1);
      const int i0 = t1 - 1 - e.z, i1 = i0 + e.w;
      if (i1 > t1 && i0 >= t0) {  // the last item starts here and goes on in other teams
        int m = 0, last = c;
        for (int j = c + 1; j < nt && s_tstart[j] < i1; ++j)
          if (s_tstart[j + 1] > s_tstart[j]) {
            ++m;
            last = j;
          }
        info.z = m;
        info.w = (last == c + m) ? 1 : 0;
      }
    }
    tinfo[c] = info;
  }
}

cudaError_t launch_sk3_plan(const int32_t* seq_lens, int batch, int hs_count, int groups, const int32_t* block_table,
                            int64_t bt_stride, int tpp_shift, int32_t* plan, Sk2Ctrl* ctrl, int calib_max, int team_force,
                            cudaStream_t stream) {
  sk3_plan_kernel<<<1, 1024, 0, stream>>>(seq_lens, batch, hs_count, groups, block_table, bt_stride, tpp_shift, plan, ctrl,
                                          calib_max, team_force);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// decode kernel
// ------------------------------------------------------------------------------------------------
template <int W, int D, int SB>
struct Sk3Smem {
  alignas(1024) uint8_t ring[W][D][SB];  // SB = bytes of one stage (K tile + V tile)
  alignas(8) uint64_t bar[W][D];         // "stage landed" mbarriers, one per ring slot
  uint32_t qstage[W][16][32];            // per warp: q fragments of the item the front prepared (cp.async, no registers)
  float m[W][kHeadsPerItem];             // merge: reference max (log2 domain) of every warp's state ...
  float l[W][kHeadsPerItem];             // ... and its row sum
  int stash[W];                          // ring slot that holds the warp's accumulators during a merge
  int word[W];                           // per team: the arrival count one thread saw (owner's look) ...
  int word2[W];                          // ... and the one its atomic returned
  long long t0;                          // clock64 at CTA start (the SM's rate is recorded at the end)
  int group;                             // chunk group this CTA claimed
};

// accumulator stash in a ring slot: [8 heads][128 d] floats, d xor-swizzled by head pair so that the
// C-fragment scatter (lane (g, t) -> head 2t | 2t + 1, d = 16 j + g) is conflict-free
__device__ __forceinline__ int sk3_stash_idx(int h, int d) { return h * kHeadDim + (d ^ (((h >> 1) & 3) << 3)); }

struct Sk3Walk {
  int b, hs, tile, n;
};

// advance a walk by `step` positions; returns true when it entered another item
__device__ __forceinline__ bool sk3_advance(const int32_t* cum, int batch, int hs_count, int step, Sk3Walk& it) {
  it.tile += step;
  bool moved = false;
  while (it.tile >= it.n) {
    it.tile -= it.n;
    moved = true;
    if (++it.hs < hs_count) continue;
    it.hs = 0;
    int c0 = __ldg(cum + it.b + 1), c1 = c0;
    do {
      ++it.b;
      c0 = c1;
      c1 = it.b < batch ? __ldg(cum + it.b + 1) : c0 + 1;
    } while (it.b < batch && c1 == c0);
    it.n = c1 - c0;
    if (it.b >= batch) {  // past the end of the stream: park
      it.tile = 0;
      it.n = 0x7fffffff;
      break;
    }
  }
  return moved;
}

// block ids of the warp's tiles t0 + step * (kbase + lane) of sequence b
__device__ __forceinline__ int sk3_load_ids(const DecodeParams& p, int b, int n_tiles, int t0, int step, int kbase, int lane) {
  const int tile = t0 + step * (kbase + lane);
  return (b < p.batch && tile < n_tiles)
             ? __ldg(p.block_table + static_cast<int64_t>(b) * p.bt_stride + (tile >> p.tpp_shift)) : 0;
}

__device__ __forceinline__ uint64_t sk3_globaltimer() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ int sk3_ld_relaxed(const int32_t* p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// running merge of one head's 4 output columns: (a, l) relative to the reference m (log2 domain)
struct Sk3Acc {
  float4 a;
  float l, m;
};
__device__ __forceinline__ void sk3_fold(Sk3Acc& s, const float4 o, float lo, float mo) {
  if (mo == -CUDART_INF_F) return;
  const float mn = fmaxf(s.m, mo);
  const float ws = (s.m == -CUDART_INF_F) ? 0.f : fast_exp2(s.m - mn);
  const float wo = fast_exp2(mo - mn);
  s.a.x = fmaf(o.x, wo, s.a.x * ws);
  s.a.y = fmaf(o.y, wo, s.a.y * ws);
  s.a.z = fmaf(o.z, wo, s.a.z * ws);
  s.a.w = fmaf(o.w, wo, s.a.w * ws);
  s.l = fmaf(lo, wo, s.l * ws);
  s.m = mn;
}

// Everything a team does at the end of a piece after the warps stashed their states: fold the T states per
// head, then write the row / leave a partial / fold the other pieces' partials (see the file header).
// Partial slots [parity][2 * NTmax][8 heads][128]: slot `team` for a team's first piece when it continues an item,
// slot NTmax + team for its last piece when it owns a cut item; arrival counters [parity][NTmax] by owner.
struct Sk3MergeArgs {
  int b, hs;          // the item
  int first, last;    // the piece starts / ends the item
  int team, tw0, T, wi;
  int own, np_first;  // tinfo.x, .y: continuation side (owner team, pieces | all-teams-in-between-non-empty << 30)
  int m_last, contig; // tinfo.z, .w: owner side
  int item_end;       // stream position behind the item
  int nt, ntmax;
  int peek;           // arrival count of the item this warp saw shortly before the piece ended (-1: did not look)
  int final_piece;    // the team's last piece: nothing is staged into the stash slots afterwards
};

struct Sk3Piece {  // one piece of a cut item as it sits in its slot: normalised row (4 columns per lane) + log2-domain LSE
  float4 o;
  float l2;
};
constexpr int kSk3FoldBatch = 2;  // partial loads in flight per lane while a cut item is folded
constexpr int kSk3FoldEarly = 2;  // pieces of an item whose partials are asked for ahead of the team barrier (incl. the own one)

template <int W, int D, int SB, bool FP8>
__device__ __noinline__ void sk3_merge(const DecodeParams& p, Sk3Smem<W, D, SB>& sm, const Sk3MergeArgs a, int warp, int lane) {
  const int parity = p.launch_id & 1;
  const int bar_id = 1 + a.tw0 / a.T, nthr = a.T * 32;
  float* const part_o = p.sk3_part_o + static_cast<int64_t>(parity) * 2 * a.ntmax * kHeadsPerItem * kHeadDim;
  float* const part_l = p.sk3_part_lse + static_cast<int64_t>(parity) * 2 * a.ntmax * kHeadsPerItem;
  int32_t* const cnt = p.sk3_cnt + parity * a.ntmax;
  const int32_t* const tstart = p.sk3_plan + sk3_plan_tstart_off();
  const bool whole = a.first && a.last;
  const bool owner = a.first && !a.last;  // cut at the end: the other pieces live in later teams
  const int pieces = owner ? a.m_last + 1 : (a.np_first & 0x3fffffff);
  // the pieces are the owner's, then the next pieces - 1 teams, when no empty team sits between them (always, but
  // for steps with fewer positions than warps)
  const bool contig = owner ? a.contig != 0 : (a.np_first >> 30) != 0;
  const int own = owner ? a.team : a.own;       // the team that holds the item's start: counter index, first slot
  const int my_idx = owner ? 0 : a.team - own;  // (contig) this piece's place in the item
  const int64_t my_slot = owner ? a.ntmax + a.team : a.team;
  const int kvh = a.hs / p.subgroups, sub = a.hs % p.subgroups;
  auto slot_of = [&](int idx) -> int64_t { return idx == 0 ? a.ntmax + own : own + idx; };  // (contig)
  auto load_piece = [&](int64_t slot, int h) {
    Sk3Piece q;
    q.o = __ldcg(reinterpret_cast<const float4*>(part_o + (slot * kHeadsPerItem + h) * kHeadDim) + lane);
    q.l2 = __ldcg(part_l + slot * kHeadsPerItem + h);
    return q;
  };
  // A piece that saw every other piece of its item arrived will finish the item: it asks for the first of their
  // partials (for this warp's first head) before the team barrier, so the loads fly while the team assembles.
  const bool saw_all = !whole && a.peek == pieces - 1;
  const bool pre = saw_all && contig && a.wi < kHeadsPerItem;
  Sk3Piece early[kSk3FoldEarly];
  if (saw_all) __threadfence();  // orders the other pieces' partials before this thread's loads
  if (pre) {
#pragma unroll
    for (int i = 0; i < kSk3FoldEarly; ++i)
      if (i < pieces && i != my_idx) early[i] = load_piece(slot_of(i), a.wi);
  }
  if (!whole && a.wi == 0 && lane == 0) sm.word[a.tw0] = a.peek;  // one thread decides for the team
  __syncwarp();
  named_bar_sync(bar_id, nthr);  // every state of the team is stashed
  // fast: every other piece had arrived when this team looked, so this team is the one that finishes the item
  const bool fast = !whole && sm.word[a.tw0] == pieces - 1;
  if (fast && !saw_all) __threadfence();
  float vs = 1.f;
  if (FP8) vs = __ldg(p.v_scale + kvh * p.scale_stride);
  // The pieces of a cut item are always folded in the same order (the owner's, then the continuations by team) and
  // form (normalised row + LSE), whoever does it and whether its own piece comes from registers or from its slot:
  // the result does not depend on which team finished last.
  auto fold_item = [&](Sk3Acc& s, int h, bool have_mine, const Sk3Piece& mine, bool have_early) {
    s.a = make_float4(0.f, 0.f, 0.f, 0.f);
    s.l = 0.f;
    s.m = -CUDART_INF_F;
    if (contig) {
      for (int base = 0; base < pieces; base += kSk3FoldBatch) {
        Sk3Piece q[kSk3FoldBatch];
#pragma unroll
        for (int i = 0; i < kSk3FoldBatch; ++i) {
          const int idx = base + i;
          if (idx >= pieces) continue;
          if (have_mine && idx == my_idx) q[i] = mine;
          else if (have_early && idx < kSk3FoldEarly) q[i] = early[i < kSk3FoldEarly ? i : 0];
          else q[i] = load_piece(slot_of(idx), h);
        }
#pragma unroll
        for (int i = 0; i < kSk3FoldBatch; ++i)
          if (base + i < pieces) sk3_fold(s, q[i].o, 1.f, q[i].l2);
      }
      return;
    }
    {
      const Sk3Piece q = (have_mine && owner) ? mine : load_piece(a.ntmax + own, h);
      sk3_fold(s, q.o, 1.f, q.l2);
    }
    for (int j = own + 1; j < a.nt && __ldg(tstart + j) < a.item_end; ++j) {
      if (!(__ldg(tstart + j + 1) > __ldg(tstart + j))) continue;
      const Sk3Piece q = (have_mine && j == a.team) ? mine : load_piece(j, h);
      sk3_fold(s, q.o, 1.f, q.l2);
    }
  };
  auto write_row = [&](const Sk3Acc& s, int h) {
    const int head_in_group = sub * kHeadsPerItem + h;
    if (head_in_group >= p.group) return;
    const int hq = kvh * p.group + head_in_group;
    const float inv = s.l > 0.f ? vs / s.l : 0.f;
    uint2 pk;
    pk.x = pack_bf16x2(s.a.x * inv, s.a.y * inv);
    pk.y = pack_bf16x2(s.a.z * inv, s.a.w * inv);
    *reinterpret_cast<uint2*>(p.out + static_cast<int64_t>(a.b) * p.out_stride_tok + static_cast<int64_t>(hq) * kHeadDim +
                              lane * 4) = pk;
    if (p.lse != nullptr && lane == 0)
      p.lse[static_cast<int64_t>(a.b) * p.num_q_heads + hq] = s.l > 0.f ? (s.m + log2f(s.l)) * kLn2 : -CUDART_INF_F;
  };
  for (int h = a.wi; h < kHeadsPerItem; h += a.T) {
    Sk3Acc s;
    s.a = make_float4(0.f, 0.f, 0.f, 0.f);
    s.l = 0.f;
    s.m = -CUDART_INF_F;
    const int col = (lane * 4) ^ (((h >> 1) & 3) << 3);
    for (int k = 0; k < a.T; ++k) {
      const float mk = sm.m[a.tw0 + k][h];
      if (mk == -CUDART_INF_F) continue;
      const float4 v = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(sm.ring[a.tw0 + k][sm.stash[a.tw0 + k]]) +
                                                        h * kHeadDim + col);
      sk3_fold(s, v, sm.l[a.tw0 + k][h], mk);
    }
    if (whole) {
      write_row(s, h);
      continue;
    }
    Sk3Piece mine;  // this team's piece as it would sit in its slot
    {
      const float inv = s.l > 0.f ? 1.f / s.l : 0.f;
      mine.o = make_float4(s.a.x * inv, s.a.y * inv, s.a.z * inv, s.a.w * inv);
      mine.l2 = s.l > 0.f ? s.m + log2f(s.l) : -CUDART_INF_F;
    }
    if (fast) {
      fold_item(s, h, true, mine, pre && h == a.wi);
      write_row(s, h);
    } else {
      *(reinterpret_cast<float4*>(part_o + (my_slot * kHeadsPerItem + h) * kHeadDim) + lane) = mine.o;
      if (lane == 0) part_l[my_slot * kHeadsPerItem + h] = mine.l2;
    }
  }
  if (whole || fast) {
    if (fast && a.wi == 0 && lane == 0) cnt[own] = 0;  // ready for the launch after the next
    if (a.final_piece) return;     // nobody refills a slot any more: the warps leave as they finish
    fence_proxy_async_smem();
    named_bar_sync(bar_id, nthr);  // the stash slots may be refilled
    return;
  }
  // a partial was written: one thread announces it (the barrier makes the team's stores visible to it, its fence to
  // the device); the last arriver of the item folds all of them
  fence_proxy_async_smem();
  named_bar_sync(bar_id, nthr);
  if (a.wi == 0 && lane == 0) {
    __threadfence();
    const int old = atomicAdd(cnt + own, 1);
    if (old == pieces - 1) __threadfence();
    sm.word2[a.tw0] = old;
  }
  named_bar_sync(bar_id, nthr);
  if (sm.word2[a.tw0] != pieces - 1) return;
  for (int h = a.wi; h < kHeadsPerItem; h += a.T) {
    Sk3Acc s;
    Sk3Piece none;
    none.o = make_float4(0.f, 0.f, 0.f, 0.f);
    none.l2 = 0.f;
    fold_item(s, h, false, none, false);
    write_row(s, h);
  }
  if (a.wi == 0 && lane == 0) cnt[own] = 0;
}

template <int W, int D, bool FP8>
__global__ void __launch_bounds__(W * 32, 1)
paged_decode_sk3_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                        const __grid_constant__ DecodeParams p) {
  constexpr int SB = FP8 ? kKvTileBytesFp8 : kKvTileBytesBf16;  // stage bytes
  constexpr int VOFF = SB / 2;                                   // V tile behind the K tile
  static_assert(SB >= kHeadsPerItem * kHeadDim * 4, "a ring slot must hold one warp's accumulators");
  extern __shared__ uint8_t smem_raw[];
  Sk3Smem<W, D, SB>& sm =
      *reinterpret_cast<Sk3Smem<W, D, SB>*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // ---- which chunk group this CTA runs: the one of the SM it sits on (the plan sized group g for SM g), claimed
  // with one atomic so that every group is run exactly once whatever the block scheduler did (two CTAs on one SM
  // one after the other, an SM id beyond the grid). A claim word holds the id of the last launch that took the group:
  // ids grow from launch to launch inside a step (the host counts the calls; the value is part of a captured graph
  // node) and the plan kernel, which runs alone once per step, zeroes the words — nothing is re-armed on the way out.
  Sk2Ctrl* const ctrl = p.sk2_ctrl;
  const int parity = p.launch_id & 1;  // counters and partial slots of consecutive launches are kept apart
  const int G = static_cast<int>(gridDim.x);
  uint32_t smid;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  const int g_spec = static_cast<int>(smid % static_cast<uint32_t>(G));
  if (threadIdx.x == 0) {
    int g = g_spec;
    if (atomicMax(&ctrl->claim3[g], p.launch_id) >= p.launch_id) {
      g = -1;
      for (int k = 1; k < G && g < 0; ++k) {
        const int gg = (g_spec + k) % G;
        if (atomicMax(&ctrl->claim3[gg], p.launch_id) < p.launch_id) g = gg;
      }
      // every word already carries this id: the caller replayed one launch id inside a step (no SM affinity then)
      if (g < 0) g = static_cast<int>(blockIdx.x);
    }
    sm.group = g;
  }
  bool waited = false;
  // (the SM's rate is measured from here: a CTA that starts while the previous kernel's stragglers still run does
  // useful work from its first instruction — it only waits before the last tile of a sequence or its first store,
  // tens of microseconds in, when the predecessor is long gone)
  if (threadIdx.x == 0) sm.t0 = clock64();
  auto wait_for_prior = [&]() {
    griddep_wait();
    waited = true;
  };
  if (p.flags & kSk2FlagWaitFirst) wait_for_prior();  // the plan was written by the kernel right before this one
  // rows of empty sequences are zeroed on the way out (no tile of theirs is in the stream): which ones is looked up
  // now, so that the exit path has no load to wait for (sequence lengths do not change during the step)
  const int eb = static_cast<int>(blockIdx.x) * W + warp;
  bool empty_rows = false;
  for (int b = eb; b < p.batch; b += G * W) empty_rows |= __ldg(p.seq_lens + b) <= 0;
  // ---- the warp's work: one round of independent loads, issued for the group the CTA will almost always get
  const int32_t* plan = p.sk3_plan;
  const int ntmax = G * W;
  const int4 hdr = __ldg(reinterpret_cast<const int4*>(plan));
  const int4 hdr2 = __ldg(reinterpret_cast<const int4*>(plan) + 1);
  const int4 hdr3 = __ldg(reinterpret_cast<const int4*>(plan) + 2);  // {teams per CTA, -, -, shares still being calibrated}
  const int T = hdr.z, nt = hdr.w, tpc = hdr3.x;
  const int wi = warp % T, tw0 = warp - wi;
  int team = g_spec * tpc + warp / T;
  int4 wd = __ldg(reinterpret_cast<const int4*>(plan + sk3_plan_wdesc_off(ntmax)) + g_spec * W + warp);
  int4 td = __ldg(reinterpret_cast<const int4*>(plan + sk3_plan_tdesc_off(ntmax)) + team);
  int4 ti = __ldg(reinterpret_cast<const int4*>(plan + sk3_plan_tinfo_off(ntmax)) + team);
  int t_begin = __ldg(plan + sk3_plan_tstart_off() + team), t_end = __ldg(plan + sk3_plan_tstart_off() + team + 1);
  int ids = __ldg(plan + sk3_plan_wids_off(ntmax) + (g_spec * W + warp) * 32 + lane);  // block ids of the first tiles
  const int32_t* cum = plan + sk3_plan_cum_off(ntmax);
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < D; ++s) mbar_init(&sm.bar[warp][s], 1);
    fence_barrier_init();
    if (warp == 0) {
      tma_prefetch_desc(&tm_k);
      tma_prefetch_desc(&tm_v);
    }
  }
  __syncthreads();
  const int group = sm.group;
  if (group != g_spec) {  // rare: this CTA runs another SM's group
    team = group * tpc + warp / T;
    wd = __ldg(reinterpret_cast<const int4*>(plan + sk3_plan_wdesc_off(ntmax)) + group * W + warp);
    ids = __ldg(plan + sk3_plan_wids_off(ntmax) + (group * W + warp) * 32 + lane);
    td = __ldg(reinterpret_cast<const int4*>(plan + sk3_plan_tdesc_off(ntmax)) + team);
    ti = __ldg(reinterpret_cast<const int4*>(plan + sk3_plan_tinfo_off(ntmax)) + team);
    t_begin = __ldg(plan + sk3_plan_tstart_off() + team);
    t_end = __ldg(plan + sk3_plan_tstart_off() + team + 1);
  }
  const int gw = group * W + warp;
  if (p.trace != nullptr && lane == 0) {  // debug: entry stamp, SM id
    p.trace[gw * 8 + 0] = sk3_globaltimer();
    p.trace[gw * 8 + 4] = smid;
  }
  // the next kernel of the chain may start its own prologue (it cannot get an SM before this CTA exits)
  griddep_launch_dependents();

  const int hs_count = p.num_kv_heads * p.subgroups;
  if (hdr.y != G || T < 1 || T > W || (W % T) != 0 || hdr2.x != p.batch || hdr2.y != hs_count || hdr2.w != kSk3PlanMagic)
    __trap();  // the plan in the workspace was made for another launch shape: fail loudly
  const bool early = (p.flags & kSk2FlagEarlyKvRead) != 0;
  const int my_first = t_begin + wi;
  const int cnt_tiles = my_first < t_end ? (t_end - my_first + T - 1) / T : 0;

  const SkLane sl = sk_make_lane(lane);
  const uint32_t ring0 = smem_u32(sm.ring[warp][0]);
  const uint32_t bar0 = smem_u32(&sm.bar[warp][0]);
  const uint64_t pol = l2_policy_evict_first();  // KV pages are streamed once per step

  // prefetch-front state
  Sk3Walk pf{wd.x, wd.y, wd.z, wd.w};
  int pf_left = cnt_tiles;
  int pf_t0 = pf.tile, pf_k = 0, pf_kbase = 0, pf_kvh = 0;  // block ids: tile pf_t0 + T * (pf_kbase + lane) of the item
  bool first_item = true;   // the block ids of the first item came with the plan
  int pf_items_ahead = 0;   // item starts prepared by the front and not yet taken by the math (0 or 1)
  bool pf_fresh = true;     // the front stands on a tile whose item it has not prepared yet
  bool q_deferred = false;  // the prepared item's q has not been loaded (griddepcontrol.wait pending)
  int it_b = 0, it_hs = 0;  // the item the front prepared last (the front itself may have moved on)
  int pslot = 0, pending = 0;
  // math state: the team's pieces in order; this warp's positions inside them
  int mb = td.x, mhs = td.y, mn = td.w;
  int item_pos0 = t_begin - td.z;  // stream position of tile 0 of the current item
  int piece_begin = t_begin;
  int mp = my_first;               // this warp's next position
  int cslot = 0;
  uint32_t cphase = 0;  // parity of the current pass over the ring
  bool have_item = false, got = false, first_landed = false;
  int peek = -1;
  uint32_t qf[8][2];
#pragma unroll
  for (int kk = 0; kk < 8; ++kk) qf[kk][0] = qf[kk][1] = 0u;
  float scale = p.scale_log2, scale_n = p.scale_log2;
  int slen = 0, slen_n = 0;
  WarpState st;
  state_reset(st);

  // q (and the K scale) of the item the front prepared last: the words of this lane's B fragments go from global
  // to the warp's staging row with cp.async (no registers held while the math is still on the previous item) — the
  // math takes them when it gets to the item.
  // bf16: qf[kk][i] = Q[head g][16 kk + 2 t + 8 i, +1]; fp8 (f16 MMAs, permuted k): Q[head g][16 kk + 4 t + 2 i, +1]
  auto load_q_next = [&]() {
    const int g = lane >> 2, t = lane & 3;
    const int kvh = it_hs / p.subgroups, sub = it_hs % p.subgroups;
    const int head_in_group = sub * kHeadsPerItem + g;
    const uint32_t* row = reinterpret_cast<const uint32_t*>(p.q + static_cast<int64_t>(it_b) * p.q_stride_tok +
                                                            static_cast<int64_t>(kvh * p.group + head_in_group) * kHeadDim);
    const uint32_t dst = smem_u32(&sm.qstage[warp][0][lane]);
    if (head_in_group < p.group) {
#pragma unroll
      for (int kk = 0; kk < 8; ++kk)
#pragma unroll
        for (int i = 0; i < 2; ++i)
          cp_async_4(dst + (kk * 2 + i) * 128, row + kk * 8 + (FP8 ? 2 * t + i : t + 4 * i));
    } else {
#pragma unroll
      for (int w = 0; w < 16; ++w) sm.qstage[warp][w][lane] = 0u;
    }
    if (FP8) scale_n = p.scale_log2 * __ldg(p.k_scale + kvh * p.scale_stride);
  };

  while (piece_begin < t_end) {
    const int item_end = item_pos0 + mn;
    const int piece_end = min(item_end, t_end);
    // a piece of a cut item looks at the item's arrival counter a few tiles before it ends (the answer is in a
    // register before the merge needs it)
    if (peek < 0 && mp + 3 * T >= piece_end) {
      if (piece_begin > item_pos0)
        peek = sk3_ld_relaxed(p.sk3_cnt + parity * ntmax + ti.x) & 0x7fffffff;   // the piece continues an item
      else if (item_end > t_end)
        peek = sk3_ld_relaxed(p.sk3_cnt + parity * ntmax + team) & 0x7fffffff;   // it owns a cut item
    }
    // ================= keep the ring full =================
    while (pending < D && pf_left > 0) {
      if (pf_fresh) {
        // first tile (of this warp) in an item: the staging row holds one prepared item, so the front never runs two item
        // starts ahead of the math
        if (pf_items_ahead == 1) break;
        it_b = pf.b;
        it_hs = pf.hs;
        pf_kvh = pf.hs / p.subgroups;
        slen_n = max(0, __ldg(p.seq_lens + pf.b));
        pf_t0 = pf.tile;
        pf_k = 0;
        pf_kbase = 0;
        if (!first_item) ids = sk3_load_ids(p, pf.b, pf.n, pf_t0, T, 0, lane);
        first_item = false;
        // q: produced by a kernel further up the stream; only with the caller's word that it was complete before this
        // kernel's predecessor started (B200ATTN_DECODE_EARLY_Q_READ) is it read ahead of griddepcontrol.wait
        if (waited || (p.flags & kSk3FlagEarlyQRead)) load_q_next(); else q_deferred = true;
        pf_items_ahead = 1;
        pf_fresh = false;
      }
      if (!waited && (!early || pf.tile + 1 == pf.n)) {
        // without the early-read contract every KV byte may still be in flight; with it only the tile that
        // holds the new token's slot
        wait_for_prior();
        if (q_deferred) {
          load_q_next();
          q_deferred = false;
        }
      }
      if (pf_k >= pf_kbase + 32) {
        pf_kbase += 32;
        ids = sk3_load_ids(p, pf.b, pf.n, pf_t0, T, pf_kbase, lane);
      }
      const int blk = __shfl_sync(0xffffffffu, ids, pf_k - pf_kbase);
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
      ++pf_k;
      pslot = (pslot + 1 == D) ? 0 : pslot + 1;
      if (--pf_left > 0) pf_fresh = sk3_advance(cum, p.batch, hs_count, T, pf);
    }
    if (q_deferred) {  // the first stages are in flight: now wait for the previous kernel and fetch q
      wait_for_prior();
      load_q_next();
      q_deferred = false;
    }
    int ds = cslot;
    if (mp < piece_end) {
      if (!have_item) {  // take the item the front prepared
        cp_async_wait_all();  // this lane's own words (nobody else reads them)
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const uint32_t raw = sm.qstage[warp][kk * 2 + i][lane];  // two bf16
            qf[kk][i] = FP8 ? pack_f16x2(__uint_as_float(raw << 16), __uint_as_float(raw & 0xffff0000u)) : raw;
          }
        scale = scale_n;
        slen = slen_n;
        pf_items_ahead = 0;
        have_item = true;
        continue;  // the front may now prepare the item after this one
      }
      // ================= math on the oldest stage =================
      mbar_wait(&sm.bar[warp][cslot], cphase);
      if (p.trace != nullptr && !first_landed && lane == 0) p.trace[gw * 8 + 1] = sk3_globaltimer();
      first_landed = true;
      {
        const int tile = mp - item_pos0;
        const int valid = (tile + 1 == mn) ? slen - tile * kTileTokens : kTileTokens;
        if (FP8)
          sk_consume_fp8<true>(ring0 + cslot * SB, sm.ring[warp][cslot], valid, qf, sl, scale, st, lane);
        else
          sk_consume<true>(ring0 + cslot * SB, sm.ring[warp][cslot], valid, qf, sl, scale, st, lane);
        if (valid < kTileTokens) fence_proxy_async_smem();  // zeroed V rows vs. the next TMA write
      }
      __syncwarp();  // every lane has read the slot before lane 0 refills it
      --pending;
      got = true;
      mp += T;
      if (++cslot == D) {
        cslot = 0;
        cphase ^= 1u;
      }
      if (mp < piece_end) continue;
    }
    // ================= the piece is done for this warp: fold the team's states =================
    {
      const int g = lane >> 2, t = lane & 3;
      if (got) {
#pragma unroll
        for (int off = 4; off <= 16; off <<= 1) {
          st.l[0] += __shfl_xor_sync(0xffffffffu, st.l[0], off);
          st.l[1] += __shfl_xor_sync(0xffffffffu, st.l[1], off);
        }
        float* stash = reinterpret_cast<float*>(sm.ring[warp][ds]);  // the slot just drained is free
        const int da = FP8 ? 2 * g : g, db = FP8 ? 2 * g + 1 : g + 8;  // fragment row -> d, see sk_consume_fp8
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          stash[sk3_stash_idx(2 * t, 16 * j + da)] = st.acc[j][0];
          stash[sk3_stash_idx(2 * t + 1, 16 * j + da)] = st.acc[j][1];
          stash[sk3_stash_idx(2 * t, 16 * j + db)] = st.acc[j][2];
          stash[sk3_stash_idx(2 * t + 1, 16 * j + db)] = st.acc[j][3];
        }
        if (g == 0) {
          sm.m[warp][2 * t] = st.m[0];
          sm.m[warp][2 * t + 1] = st.m[1];
          sm.l[warp][2 * t] = st.l[0];
          sm.l[warp][2 * t + 1] = st.l[1];
        }
        if (lane == 0) sm.stash[warp] = ds;
      } else if (lane < kHeadsPerItem) {
        sm.m[warp][lane] = -CUDART_INF_F;
        sm.l[warp][lane] = 0.f;
      }
      if (!waited) wait_for_prior();  // the merge writes global memory
      Sk3MergeArgs a;
      a.b = mb;
      a.hs = mhs;
      a.first = piece_begin == item_pos0;
      a.last = piece_end == item_end;
      a.team = team;
      a.tw0 = tw0;
      a.T = T;
      a.wi = wi;
      a.own = ti.x;
      a.np_first = ti.y;
      a.m_last = ti.z;
      a.contig = ti.w;
      a.item_end = item_end;
      a.nt = nt;
      a.ntmax = ntmax;
      a.peek = peek;
      a.final_piece = piece_end == t_end;
      sk3_merge<W, D, SB, FP8>(p, sm, a, warp, lane);
      state_reset(st);
      got = false;
      have_item = false;  // (nothing of the item is kept across the merge: the next piece takes its own q)
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) qf[kk][0] = qf[kk][1] = 0u;
      scale = p.scale_log2;
      slen = 0;
      peek = -1;
    }
    piece_begin = piece_end;
    if (piece_end == item_end && piece_begin < t_end) {  // next item of the team's range
      item_pos0 = item_end;
      if (++mhs == hs_count) {
        mhs = 0;
        int c0 = __ldg(cum + mb + 1), c1 = c0;
        do {
          ++mb;
          c0 = c1;
          c1 = mb < p.batch ? __ldg(cum + mb + 1) : c0 + 1;
        } while (mb < p.batch && c1 == c0);
        mn = c1 - c0;
      }
    }
  }

  // every warp waits before it exits, so a completed grid implies its predecessors completed
  if (!waited) wait_for_prior();
  if (p.trace != nullptr && lane == 0) {  // debug: end of this warp's work
    const uint64_t t_end_ns = sk3_globaltimer();
    if (!first_landed) p.trace[gw * 8 + 1] = t_end_ns;
    p.trace[gw * 8 + 2] = t_end_ns;
    p.trace[gw * 8 + 3] = t_end_ns;
  }
  // rows of empty sequences: zero, LSE = -inf
  if (empty_rows) {
    for (int b = eb; b < p.batch; b += G * W) {
      if (__ldg(p.seq_lens + b) > 0) continue;
      for (int i = lane; i < p.num_q_heads * (kHeadDim / 4); i += 32)
        *reinterpret_cast<uint2*>(p.out + static_cast<int64_t>(b) * p.out_stride_tok + i * 4) = make_uint2(0u, 0u);
      if (p.lse != nullptr)
        for (int h = lane; h < p.num_q_heads; h += 32) p.lse[static_cast<int64_t>(b) * p.num_q_heads + h] = -CUDART_INF_F;
    }
  }
  if (hdr3.w == 0) return;  // the SM shares are frozen: nothing to record, every warp leaves on its own
  __syncthreads();
  if (threadIdx.x == 0 && smid < static_cast<uint32_t>(kSk2MaxGroups)) {
    // what this SM achieved (the plan kernel turns it into the SM's share of the next steps' streams), summed over
    // the launches since the last plan (every layer of a step): one launch is a noisy sample
    const int32_t* tstart = plan + sk3_plan_tstart_off();
    const long long cyc = clock64() - sm.t0;
    if (cyc > 0 && cyc < 0x10000000ll && ctrl->stat_cycles[smid] < 0x30000000) {
      atomicAdd(&ctrl->stat_tiles[smid], __ldg(tstart + (group + 1) * tpc) - __ldg(tstart + group * tpc));
      atomicAdd(&ctrl->stat_cycles[smid], static_cast<int>(cyc));
    }
  }
}

// Warps per CTA: 12 = 3 per SM sub-partition, the most that still leaves each thread the ~168 registers the
// streaming loop needs; ring stages per warp: 2 x 8 KiB (bf16), 4 x 4 KiB (fp8).
constexpr int kSk3Depth = 2;

size_t decode_sk3_smem_bytes(bool fp8) {
  return (fp8 ? sizeof(Sk3Smem<kSk3Warps, 2 * kSk3Depth, kKvTileBytesFp8>)
              : sizeof(Sk3Smem<kSk3Warps, kSk3Depth, kKvTileBytesBf16>)) + 1024;
}

template <typename K>
static cudaError_t sk3_launch(K kern, int* configured_dev, size_t smem, int threads, const CUtensorMap& tm_k,
                              const CUtensorMap& tm_v, const DecodeParams& p, int grid, bool pdl, cudaStream_t stream) {
  int dev = 0;
  if (cudaError_t e = cudaGetDevice(&dev)) return e;
  if (dev != *configured_dev) {
    if (cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem)))
      return e;
    *configured_dev = dev;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(static_cast<unsigned>(threads));
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, tm_k, tm_v, p);
}

cudaError_t launch_decode_sk3(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const DecodeParams& p, bool fp8,
                              int grid, bool pdl, cudaStream_t stream) {
  static thread_local int dev[2] = {-1, -1};
  if (fp8)
    return sk3_launch(paged_decode_sk3_kernel<kSk3Warps, 2 * kSk3Depth, true>, &dev[1], decode_sk3_smem_bytes(true),
                      kSk3Warps * 32, tm_k, tm_v, p, grid, pdl, stream);
  return sk3_launch(paged_decode_sk3_kernel<kSk3Warps, kSk3Depth, false>, &dev[0], decode_sk3_smem_bytes(false),
                    kSk3Warps * 32, tm_k, tm_v, p, grid, pdl, stream);
}

}  // namespace b200
