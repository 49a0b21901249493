// Split-K paged decode attention for sm_100a, bf16 KV cache — persistent "stream" variant.
//
// One CTA per SM, W warps. Every warp owns a private ring of D stages (one 16-token K tile plus
// its V tile = 8 KiB per stage) that it fills itself with 16-byte cp.async and drains with the
// tensor-core math of decode_common.cuh, so the steady state needs no cross-warp synchronisation
// at all. A warp's stages form ONE continuous stream over all work items of its CTA: the
// prefetch front runs up to D stages ahead of the math and crosses work-item boundaries, so the
// per-item prologue (first page latency) and epilogue (cross-warp merge + store) no longer drain
// the HBM pipe. Measured ceiling of this staging method on B200: ~7.1 TB/s of random 8 KiB
// pages (scripts/probe/stream_probe.cu), against ~3.4 TB/s for 2 KiB TMA boxes.
//
// Item boundary: the W warps merge their partial (m, l, O) through shared memory. The merge
// buffer is not extra memory: each warp parks its O in the ring slot it has just drained (the
// prefetch of the following stage is issued after the merge), while its other D-1 slots keep
// streaming the next item's pages.
//
// Replaces the decode branch of the reference's stock backends
// (vllm/v1/attention/backends/flashinfer.py:1694-1822, flash_attn.py:758-819).
#include "kernels.h"

namespace b200 {

constexpr int kStreamWarps = 8;
constexpr int kStreamDepth = 3;

template <int W, int D>
struct StreamSmem {
  alignas(1024) uint8_t ring[W][D][kKvTileBytesBf16];
  float m[W][kHeadsPerItem];
  float l[W][kHeadsPerItem];
  int slot[W];
};

// What a warp knows about one work item of its CTA.
struct StreamItem {
  WorkItem w;
  const int32_t* bt;  // block-table row
  int nst;            // stages (16-token tiles) of THIS warp in the item
  bool valid;         // item index < total_items
  bool active;        // produces output
};

template <int W>
__device__ __forceinline__ void stream_item_init(const DecodeParams& p, int item, int seq_len, int warp,
                                                 StreamItem& s) {
  s.valid = item < p.total_items;
  if (!s.valid) {
    s.nst = 0;
    s.active = false;
    s.bt = p.block_table;
    s.w.b = s.w.kvh = s.w.split = s.w.sub = 0;
    s.w.seq_len = s.w.tile_begin = s.w.tile_end = 0;
    s.w.actual_splits = 1;
    return;
  }
  item_coords(p, item, s.w);
  item_finish(p, seq_len, s.w);
  s.bt = p.block_table + static_cast<int64_t>(s.w.b) * p.bt_stride;
  const int ntiles = s.w.tile_end - s.w.tile_begin;
  s.nst = ntiles > warp ? (ntiles - warp + W - 1) / W : 0;
  s.active = item_active(s.w);
}

// Sequence length of item `item` (0 when past the end): issued one item early so that the
// dependent index math never waits on it.
__device__ __forceinline__ int stream_seq_len(const DecodeParams& p, int item) {
  if (item >= p.total_items) return 0;
  WorkItem w;
  item_coords(p, item, w);
  return __ldg(p.seq_lens + w.b);
}

// Block ids of this warp's stages [32*win, 32*win + 32) of the item, one per lane.
template <int W>
__device__ __forceinline__ int stream_load_ids(const DecodeParams& p, const StreamItem& s, int warp, int lane,
                                               int win) {
  const int tile = s.w.tile_begin + warp + (win * 32 + lane) * W;
  return (s.valid && tile < s.w.tile_end) ? __ldg(s.bt + tile / p.tiles_per_page) : 0;
}

// Stage `k` (warp-local index) of the item -> ring slot `dst`, SWIZZLE_128B layout.
template <int W>
__device__ __forceinline__ void stream_issue(const DecodeParams& p, const StreamItem& s, int ids, int k, int warp,
                                             int lane, uint8_t* dst) {
  const int tile = s.w.tile_begin + warp + k * W;
  const int64_t blk = __shfl_sync(0xffffffffu, ids, k & 31);
  const int slot0 = (tile % p.tiles_per_page) * kTileTokens;
  const int c = lane & 15;     // 16-byte chunk of the 256-byte row
  const int rpar = lane >> 4;  // row parity handled by this half-warp
  const int64_t base_b =
      (blk * p.cs_block + static_cast<int64_t>(s.w.kvh) * p.cs_head + static_cast<int64_t>(slot0 + rpar) * p.cs_slot) * 2 +
      c * 16;
  const int64_t row2_b = p.cs_slot * 4;  // two rows, bf16
  const uint8_t* ks = p.k_cache + base_b;
  const uint8_t* vs = p.v_cache + base_b;
  uint8_t* kt = dst + (c >> 3) * kHalfBytes;
  uint8_t* vt = kt + kTileBytesBf16;
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int r = 2 * it + rpar;
    const uint32_t off = swz_off(r, c & 7);
    cp_async_16(kt + off, ks + it * row2_b);
    cp_async_16(vt + off, vs + it * row2_b);
  }
}

__device__ __forceinline__ void cp_async_wait_dyn(int allowed_pending) {
  switch (allowed_pending) {
    case 0: cp_async_wait<0>(); break;
    case 1: cp_async_wait<1>(); break;
    case 2: cp_async_wait<2>(); break;
    case 3: cp_async_wait<3>(); break;
    default: cp_async_wait<4>(); break;
  }
}

template <int W, int D>
__global__ void __launch_bounds__(W * 32, 1) paged_decode_stream_kernel(const DecodeParams p) {
  extern __shared__ uint8_t smem_raw[];
  StreamSmem<W, D>& sm =
      *reinterpret_cast<StreamSmem<W, D>*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int G = gridDim.x;
  const LaneOffsets lo = make_lane_offsets(lane);

  // ---- items: cur (math), nxt (prefetch may run into it), and the sequence length after that
  int item = blockIdx.x;
  StreamItem cur, nxt;
  {
    const int s0 = stream_seq_len(p, item), s1 = stream_seq_len(p, item + G);
    stream_item_init<W>(p, item, s0, warp, cur);
    stream_item_init<W>(p, item + G, s1, warp, nxt);
  }
  int seq_nn = stream_seq_len(p, item + 2 * G);
  uint32_t qf[8][2], qf_n[8][2];
  load_q_frags(p, cur.w, lane, qf);
  load_q_frags(p, nxt.w, lane, qf_n);
  int ids_c = stream_load_ids<W>(p, cur, warp, lane, 0);
  int ids_n = stream_load_ids<W>(p, nxt, warp, lane, 0);

  bool pf_in_nxt = false;  // prefetch front has moved on to `nxt`
  int pk = 0;              // next warp-local stage the prefetch front issues (within its item)
  int pslot = 0;           // ring slot it goes to
  int pending = 0;         // stages issued and not yet consumed
  int ck = 0;              // next warp-local stage the math consumes (within cur)
  int cslot = 0;
  WarpState st;
  state_reset(st);

  while (cur.valid) {
    if (ck == cur.nst) {
      // ================= item boundary: merge the W warps, store, advance =================
      if (cur.active) {
        const int g = lane >> 2, t = lane & 3;
#pragma unroll
        for (int off = 4; off <= 16; off <<= 1) {
          st.l[0] += __shfl_xor_sync(0xffffffffu, st.l[0], off);
          st.l[1] += __shfl_xor_sync(0xffffffffu, st.l[1], off);
        }
        // the slot drained last is free (pending <= D-1 here): park O there
        const int fslot = (cslot + D - 1) % D;
        if (g == 0) {
          sm.m[warp][2 * t] = st.m[0];
          sm.m[warp][2 * t + 1] = st.m[1];
          sm.l[warp][2 * t] = st.l[0];
          sm.l[warp][2 * t + 1] = st.l[1];
        }
        if (lane == 0) sm.slot[warp] = fslot;
        if (cur.nst > 0) {
          float* o = reinterpret_cast<float*>(sm.ring[warp][fslot]);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            o[(2 * t) * kMergeRow + 16 * j + g] = st.acc[j][0];
            o[(2 * t + 1) * kMergeRow + 16 * j + g] = st.acc[j][1];
            o[(2 * t) * kMergeRow + 16 * j + g + 8] = st.acc[j][2];
            o[(2 * t + 1) * kMergeRow + 16 * j + g + 8] = st.acc[j][3];
          }
        }
        __syncthreads();
        const WorkItem& w = cur.w;
        const bool direct = (w.actual_splits == 1);
        for (int idx = threadIdx.x; idx < kHeadsPerItem * (kHeadDim / 4); idx += W * 32) {
          const int head = idx / (kHeadDim / 4);
          const int d0 = (idx % (kHeadDim / 4)) * 4;
          const int head_in_group = w.sub * kHeadsPerItem + head;
          if (head_in_group >= p.group) continue;
          const int h = w.kvh * p.group + head_in_group;
          float M = -CUDART_INF_F;
#pragma unroll
          for (int ww = 0; ww < W; ++ww) M = fmaxf(M, sm.m[ww][head]);
          float denom = 0.f, v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
          if (M != -CUDART_INF_F) {
#pragma unroll
            for (int ww = 0; ww < W; ++ww) {
              const float mw = sm.m[ww][head];
              if (mw == -CUDART_INF_F) continue;  // warp had no tokens (its slot holds no O)
              const float wgt = fast_exp2(mw - M);
              denom = fmaf(sm.l[ww][head], wgt, denom);
              const float* o = reinterpret_cast<const float*>(sm.ring[ww][sm.slot[ww]]);
              const float4 ov = *reinterpret_cast<const float4*>(o + head * kMergeRow + d0);
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
            uint2 pk2;
            pk2.x = pack_bf16x2(v0, v1);
            pk2.y = pack_bf16x2(v2, v3);
            *reinterpret_cast<uint2*>(dst) = pk2;
            if (p.lse != nullptr && d0 == 0) p.lse[static_cast<int64_t>(w.b) * p.num_q_heads + h] = lse_nat;
          } else {
            const int64_t row = (static_cast<int64_t>(w.b) * p.num_q_heads + h) * p.num_splits + w.split;
            *reinterpret_cast<float4*>(p.part_o + row * kHeadDim + d0) = make_float4(v0, v1, v2, v3);
            if (d0 == 0) p.part_lse[row] = lse_nat;
          }
        }
        __syncthreads();  // parked O has been read: the slots may be refilled
      }
      // advance: cur <- nxt, decode the item after it, prefetch the sequence length after that
      item += G;
      cur = nxt;
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        qf[kk][0] = qf_n[kk][0];
        qf[kk][1] = qf_n[kk][1];
      }
      ids_c = ids_n;
      if (!pf_in_nxt) pk = 0;  // the front was still at the end of the old item
      pf_in_nxt = false;
      stream_item_init<W>(p, item + G, seq_nn, warp, nxt);
      seq_nn = stream_seq_len(p, item + 2 * G);
      load_q_frags(p, nxt.w, lane, qf_n);
      ids_n = stream_load_ids<W>(p, nxt, warp, lane, 0);
      ck = 0;
      state_reset(st);
      continue;
    }

    // ================= keep the ring full =================
    while (pending < D) {
      if (!pf_in_nxt) {
        if (pk >= cur.nst) {
          pf_in_nxt = true;
          pk = 0;
          continue;
        }
        if ((pk & 31) == 0 && pk > 0) ids_c = stream_load_ids<W>(p, cur, warp, lane, pk >> 5);
        stream_issue<W>(p, cur, ids_c, pk, warp, lane, sm.ring[warp][pslot]);
      } else {
        if (pk >= nxt.nst) break;  // the front never runs past `nxt`
        if ((pk & 31) == 0 && pk > 0) ids_n = stream_load_ids<W>(p, nxt, warp, lane, pk >> 5);
        stream_issue<W>(p, nxt, ids_n, pk, warp, lane, sm.ring[warp][pslot]);
      }
      cp_async_commit();
      ++pending;
      ++pk;
      pslot = (pslot + 1 == D) ? 0 : pslot + 1;
    }

    // ================= math on the oldest stage =================
    {
      cp_async_wait_dyn(pending - 1);
      __syncwarp();
      const int tile = cur.w.tile_begin + warp + ck * W;
      const int valid = min(kTileTokens, cur.w.seq_len - tile * kTileTokens);
      consume_stage_bf16<1>(sm.ring[warp][cslot], valid, valid < kTileTokens, qf, lo, p.scale_log2, st, lane);
      __syncwarp();
      --pending;
      ++ck;
      cslot = (cslot + 1 == D) ? 0 : cslot + 1;
    }
  }
  cp_async_wait<0>();
}

size_t decode_stream_smem_bytes() { return sizeof(StreamSmem<kStreamWarps, kStreamDepth>) + 1024; }

cudaError_t launch_decode_stream(const DecodeParams& p, int grid, cudaStream_t stream) {
  auto kern = paged_decode_stream_kernel<kStreamWarps, kStreamDepth>;
  static thread_local int configured_dev = -1;
  const size_t smem = decode_stream_smem_bytes();
  int dev = 0;
  if (cudaError_t e = cudaGetDevice(&dev)) return e;
  if (dev != configured_dev) {
    if (cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem)))
      return e;
    configured_dev = dev;
  }
  kern<<<grid, kStreamWarps * 32, smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace b200
