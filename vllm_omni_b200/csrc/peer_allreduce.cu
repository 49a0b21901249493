// Peer-memory all-reduce (sum, bf16) over NVLink 5 / NVSwitch for the o_proj row-parallel sum that
// follows the attention path (north_star: "the downstream o_proj all-reduce"; the reference reaches it
// through vLLM's RowParallelLinear -> tensor_model_parallel_all_reduce, vllm/distributed/
// communication_op.py:12, which on NVLink boxes dispatches to custom_all_reduce.cuh or NCCL).
//
// One process per GPU. Every rank owns one peer-mapped allocation (torch symmetric memory or
// cudaIpc, the host layer's business) carved into [flags | input | scratch]; the kernel gets the
// per-rank base pointers BY VALUE (<= 8 ranks) and never touches the host. Three schedules:
//
//   ONESHOT   every rank loads the 16-byte vector of all ranks and sums it in rank order in fp32.
//             1 NVLink read per element per peer, 2 flag barriers. Best below ~256 KiB x world.
//   TWOSHOT   reduce-scatter + all-gather inside ONE launch: rank r sums slice r of every peer into its
//             scratch, a flag barrier publishes it, then every rank copies the other slices. (world-1)/world
//             x 2 of the bytes cross NVLink instead of (world-1) x.
//   PUSH      barrier-free ("Lamport") one-shot for latency-bound sizes: every rank WRITES its vectors into a
//             slot of each peer's receive buffer and the receivers spin on the data itself — the buffers are
//             pre-filled with a sentinel (bf16 -0.0, which senders canonicalise to +0.0, so no sum changes) and
//             a 16-byte vector is complete when none of its halves is the sentinel. One posted NVLink write of
//             latency instead of two flag round trips; three receive buffers rotate (step k uses k % 3 and
//             re-arms the one of step k-1), so nothing is ever reset by the host. The input needs no peer mapping.
//   PUSH2     the same idea for bandwidth-sized messages on 4/8 ranks: barrier-free reduce-scatter + all-gather
//             (copies of slice p are posted to rank p, which adds them in rank order and posts the sum to all).
//   MULTIMEM  NVLS: `multimem.ld_reduce` on the multicast address lets the NVSwitch add the replicas
//             (fp32 accumulate), `multimem.st` broadcasts the sum; each rank handles slice r. Needs a
//             multicast mapping of the allocation.
//
// Flag protocol (CUDA-graph safe, nothing is reset between launches): flag words are monotonically
// increasing launch-local epochs. Block b of rank r owns epoch counter `epoch[b]` in its own flag
// region; a barrier = bump the epoch, store it into slot [b][r] of every peer's flag region with
// a relaxed system-scope store, spin with relaxed loads on the `world` slots of the own region until each is >= the
// epoch (wrap-safe signed compare). A peer can be at most one barrier ahead, so >= is sufficient. Same-index
// blocks pair up across ranks, and both TWOSHOT phases use the same (block, thread) -> element mapping, so a
// per-block barrier orders exactly the data that block consumes. The sum order is rank 0..world-1 in fp32
// with one final bf16 rounding, so ONESHOT/TWOSHOT results are bit-identical on every rank and equal to the
// sequential fp32 oracle. A spin that exceeds `timeout_ns` records the stuck (block, peer) in the flag region
// and traps: a dead peer fails the launch loudly instead of hanging the box.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels.h"

namespace b200 {

namespace {

__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// flag region layout (uint32 words): [kPeerMaxBlocks][8] arrival slots | [kPeerMaxBlocks] epochs | 4 error words
__device__ __forceinline__ uint32_t* slot_of(uint32_t* flags, int block, int from_rank) {
  return flags + block * kPeerMaxRanks + from_rank;
}
__device__ __forceinline__ uint32_t* epoch_of(uint32_t* flags, int block) {
  return flags + kPeerMaxBlocks * kPeerMaxRanks + block;
}
__device__ __forceinline__ uint32_t* error_of(uint32_t* flags) {
  return flags + kPeerMaxBlocks * kPeerMaxRanks + kPeerMaxBlocks;
}
// PUSH schedule state: {step, blocks done, vectors of the previous step, -}
__device__ __forceinline__ uint32_t* push_state_of(uint32_t* flags) { return error_of(flags) + 4; }

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Block-wide barrier with the same-index block of every peer.
//   fence = true : data written by this block before the barrier is visible to the peers after it
//                  (st.release.sys flag, one ld.acquire.sys once the flag is seen; bar.sync makes both
//                  cumulative over the block). System-scope release/acquire drains to NVLink and costs
//                  microseconds (measured: +3-4 us per barrier, two explicit fence.sys even more), so it is
//                  paid only where a peer consumes what this launch produced.
//   fence = false: control only, relaxed flag traffic. Enough (a) at kernel start, where the data the peers
//                  read was completed by an EARLIER kernel of their stream (visible at kernel completion), and
//                  (b) for "I have finished reading your buffer", where every load's value was consumed
//                  before the bar.sync below.
__device__ __forceinline__ void peer_barrier(const PeerCtl& a, uint32_t& epoch, bool fence) {
  __syncthreads();
  epoch += 1;
  const int t = threadIdx.x;
  if (t < a.world) {
    if (fence)
      st_release_sys(slot_of(a.flags[t], blockIdx.x, a.rank), epoch);
    else
      st_relaxed_sys(slot_of(a.flags[t], blockIdx.x, a.rank), epoch);
    const uint32_t* mine = slot_of(a.flags[a.rank], blockIdx.x, t);
    const uint64_t t0 = globaltimer_ns();
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_relaxed_sys(mine) - epoch) < 0) {
      if ((++spins & 0x3ff) == 0 && globaltimer_ns() - t0 > a.timeout_ns) {
        uint32_t* e = error_of(a.flags[a.rank]);
        e[0] = 1u; e[1] = blockIdx.x; e[2] = static_cast<uint32_t>(t); e[3] = epoch;
        __threadfence_system();
        if (!a.soft_timeout) __trap();
        break;  // test mode: give up on this wait, the host reads the error words
      }
    }
    if (fence) (void)ld_acquire_sys(mine);   // the spin saw the flag relaxed: one acquire load orders what follows
  }
  __syncthreads();
}

// Programmatic dependent launch: these kernels are launched with programmaticStreamSerialization, so their
// CTAs are scheduled while the previous kernel of the stream drains; nothing may be touched before the wait.
// Releasing the dependents first is safe because every kernel of this library waits the same way.
__device__ __forceinline__ void pdl_prologue() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ void acc_bf16x8(float (&acc)[8], const uint4& v) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(h[i]);
    acc[2 * i] += f.x;
    acc[2 * i + 1] += f.y;
  }
}
__device__ __forceinline__ uint4 pack_bf16x8(const float (&acc)[8]) {
  uint4 v;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]);
  return v;
}

// sum of vector `idx` over all ranks, rank order, fp32, one rounding
template <int WORLD>
__device__ __forceinline__ uint4 reduce_vec(const PeerArgs& a, int64_t idx) {
  uint4 v[WORLD];
#pragma unroll
  for (int r = 0; r < WORLD; ++r) v[r] = ld_peer_v4(reinterpret_cast<const uint4*>(a.in[r]) + idx);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int r = 0; r < WORLD; ++r) acc_bf16x8(acc, v[r]);
  return pack_bf16x8(acc);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// kernels. n_vec = number of 16-byte vectors (8 bf16 each); the host pads numel to a multiple of 8.
// ---------------------------------------------------------------------------------------------
template <int WORLD>
__global__ void __launch_bounds__(kPeerThreads) peer_allreduce_oneshot_kernel(PeerArgs a, int64_t n_vec) {
  pdl_prologue();
  uint32_t epoch = *epoch_of(a.ctl.flags[a.ctl.rank], blockIdx.x);
  peer_barrier(a.ctl, epoch, /*fence=*/false);  // every peer's producer kernel has finished (stream order + kernel boundary)
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kPeerThreads;
  uint4* out = reinterpret_cast<uint4*>(a.out);
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * kPeerThreads + threadIdx.x; i < n_vec; i += stride)
    out[i] = reduce_vec<WORLD>(a, i);
  peer_barrier(a.ctl, epoch, /*fence=*/false);  // nobody still reads my input when this launch retires
  if (threadIdx.x == 0) *epoch_of(a.ctl.flags[a.ctl.rank], blockIdx.x) = epoch;
}

template <int WORLD>
__global__ void __launch_bounds__(kPeerThreads) peer_allreduce_twoshot_kernel(PeerArgs a, int64_t n_vec) {
  pdl_prologue();
  uint32_t epoch = *epoch_of(a.ctl.flags[a.ctl.rank], blockIdx.x);
  const int64_t part = n_vec / WORLD;                 // slice r = [r*part, (r+1)*part), the last one takes the tail
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * kPeerThreads + threadIdx.x;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kPeerThreads;
  const int64_t my_start = a.ctl.rank * part;
  const int64_t my_len = (a.ctl.rank == WORLD - 1) ? n_vec - my_start : part;
  peer_barrier(a.ctl, epoch, false);
  uint4* scratch = reinterpret_cast<uint4*>(a.scratch[a.ctl.rank]);
  for (int64_t j = tid; j < my_len; j += stride) scratch[j] = reduce_vec<WORLD>(a, my_start + j);
  peer_barrier(a.ctl, epoch, /*fence=*/true);             // reduced slices are published
  uint4* out = reinterpret_cast<uint4*>(a.out);
  const int64_t longest = n_vec - (WORLD - 1) * part;
  for (int64_t j = tid; j < longest; j += stride) {
#pragma unroll
    for (int i = 0; i < WORLD; ++i) {
      const int r = (a.ctl.rank + i) % WORLD;             // start with the local slice, spread the peers
      const int64_t len = (r == WORLD - 1) ? longest : part;
      if (j < len) out[r * part + j] = ld_peer_v4(reinterpret_cast<const uint4*>(a.scratch[r]) + j);
    }
  }
  peer_barrier(a.ctl, epoch, false);                      // scratch and input may be overwritten by the next launch
  if (threadIdx.x == 0) *epoch_of(a.ctl.flags[a.ctl.rank], blockIdx.x) = epoch;
}

// ---- PUSH ("Lamport") one-shot --------------------------------------------------------------
constexpr uint32_t kSentinelWord = 0x80008000u;  // two bf16 -0.0
__device__ __forceinline__ uint32_t canon_word(uint32_t w) {  // -0.0 -> +0.0 in either half
  if ((w & 0xffffu) == 0x8000u) w &= 0xffff0000u;
  if ((w >> 16) == 0x8000u) w &= 0x0000ffffu;
  return w;
}
__device__ __forceinline__ bool word_pending(uint32_t w) { return (w & 0xffffu) == 0x8000u || (w >> 16) == 0x8000u; }
__device__ __forceinline__ bool vec_pending(const uint4& v) {
  return word_pending(v.x) || word_pending(v.y) || word_pending(v.z) || word_pending(v.w);
}
__device__ __forceinline__ void st_peer_v4(void* p, const uint4& x) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(x.x), "r"(x.y), "r"(x.z), "r"(x.w)
               : "memory");
}

// a.push[r] = rank r's receive region: 3 rotating buffers of WORLD * slot_vec vectors each. Layout inside a buffer:
//   PUSH  (kind 1): [WORLD slots][slot_vec]                       slot s = what rank s pushed (whole message)
//   PUSH2 (kind 2): [A: WORLD slots][slot_vec/2] [B: WORLD slots][slot_vec/2]
//                   A[s] = rank s's copy of MY slice (to be reduced here), B[s] = slice s already reduced by rank s
// a.in[rank] is this rank's input (any device memory); a.out the local result.
// State words (push_state_of): {step, blocks done, vectors of the previous step, kind of the previous step}.
__device__ __forceinline__ int64_t slice_len(int64_t n_vec, int world, int r) {
  const int64_t part = n_vec / world;
  return r == world - 1 ? n_vec - part * (world - 1) : part;
}

// spin until the 16-byte vector at `src` (local memory a peer writes) holds no sentinel half
__device__ __forceinline__ uint4 poll_vec(const PeerArgs& a, const uint4* src, uint64_t t0, int from, uint32_t step) {
  uint32_t spins = 0;
  uint4 v;
  while (true) {
    v = ld_peer_v4(src);
    if (!vec_pending(v)) break;
    if ((++spins & 0x3ff) == 0 && globaltimer_ns() - t0 > a.ctl.timeout_ns) {
      uint32_t* e = error_of(a.ctl.flags[a.ctl.rank]);
      e[0] = 2u; e[1] = blockIdx.x; e[2] = static_cast<uint32_t>(from); e[3] = step;
      __threadfence_system();
      if (!a.ctl.soft_timeout) __trap();
      break;
    }
  }
  return v;
}

// re-arm what the previous step received in buffer `prev` (all of it was consumed before that launch retired)
template <int WORLD>
__device__ __forceinline__ void push_rearm(uint4* mine, int prev, uint32_t kind, int64_t n_prev, int64_t slot_vec,
                                           int rank, int64_t tid, int64_t stride) {
  const uint4 sent = make_uint4(kSentinelWord, kSentinelWord, kSentinelWord, kSentinelWord);
  uint4* buf = mine + static_cast<int64_t>(prev) * WORLD * slot_vec;
  if (kind == 1u) {
    for (int64_t i = tid; i < n_prev; i += stride) {
#pragma unroll
      for (int r = 0; r < WORLD; ++r)
        if (r != rank) buf[r * slot_vec + i] = sent;
    }
  } else if (kind == 2u) {
    const int64_t cap = slot_vec / 2, part = n_prev / WORLD;
    const int64_t my_len = slice_len(n_prev, WORLD, rank), longest = slice_len(n_prev, WORLD, WORLD - 1);
    uint4* bufB = buf + WORLD * cap;
    for (int64_t j = tid; j < longest; j += stride) {
#pragma unroll
      for (int r = 0; r < WORLD; ++r) {
        if (r == rank) continue;
        if (j < my_len) buf[r * cap + j] = sent;
        if (j < (r == WORLD - 1 ? longest : part)) bufB[r * cap + j] = sent;
      }
    }
  }
}

// the last block to finish advances the step (every block has read the state by then)
__device__ __forceinline__ void push_advance(uint32_t* state, uint32_t step, int64_t n_vec, uint32_t kind) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t done = atomicAdd(state + 1, 1u);
    if (done == gridDim.x - 1) {
      state[1] = 0u;
      state[2] = static_cast<uint32_t>(n_vec);
      state[3] = kind;
      state[0] = step + 1;
      __threadfence();
    }
  }
}

__device__ __forceinline__ uint4 canon_vec(uint4 v) {
  v.x = canon_word(v.x); v.y = canon_word(v.y); v.z = canon_word(v.z); v.w = canon_word(v.w);
  return v;
}

template <int WORLD>
__global__ void __launch_bounds__(kPeerThreads) peer_allreduce_push_kernel(PeerArgs a, int64_t n_vec, int64_t slot_vec) {
  pdl_prologue();
  const int rank = a.ctl.rank;
  uint32_t* state = push_state_of(a.ctl.flags[rank]);
  const uint32_t step = state[0], prev_kind = state[3];
  const int64_t n_prev = state[2];
  const int cur = step % 3, prev = (step + 2) % 3;
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * kPeerThreads + threadIdx.x;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kPeerThreads;
  const uint4* in = reinterpret_cast<const uint4*>(a.in[rank]);
  uint4* mine = reinterpret_cast<uint4*>(a.push[rank]);
  // 1. push my vectors into slot `rank` of every peer's current buffer (posted writes, no round trip)
  for (int64_t i = tid; i < n_vec; i += stride) {
    const uint4 v = canon_vec(in[i]);
#pragma unroll
    for (int d = 1; d < WORLD; ++d) {
      const int p = (rank + d) % WORLD;
      st_peer_v4(reinterpret_cast<uint4*>(a.push[p]) + (static_cast<int64_t>(cur) * WORLD + rank) * slot_vec + i, v);
    }
  }
  // 2. re-arm the buffer of the previous step
  push_rearm<WORLD>(mine, prev, prev_kind, n_prev, slot_vec, rank, tid, stride);
  // 3. wait for the peers' vectors to land in my current buffer, add in rank order
  const uint64_t t0 = globaltimer_ns();
  for (int64_t i = tid; i < n_vec; i += stride) {
    uint4 v[WORLD];
    v[rank] = in[i];
#pragma unroll
    for (int r = 0; r < WORLD; ++r)
      if (r != rank) v[r] = poll_vec(a, mine + (static_cast<int64_t>(cur) * WORLD + r) * slot_vec + i, t0, r, step);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < WORLD; ++r) acc_bf16x8(acc, v[r]);
    reinterpret_cast<uint4*>(a.out)[i] = pack_bf16x8(acc);
  }
  push_advance(state, step, n_vec, 1u);
}

// PUSH2: barrier-free reduce-scatter + all-gather. Rank r owns slice r: every rank posts its copy of slice p
// to rank p (A slots), rank p adds the copies in rank order and posts the sum to everybody (B slots). Two posted
// writes of latency, 2 (world-1)/world of the bytes per rank on the wire; no flags, no fences.
template <int WORLD>
__global__ void __launch_bounds__(kPeerThreads) peer_allreduce_push2_kernel(PeerArgs a, int64_t n_vec, int64_t slot_vec) {
  pdl_prologue();
  const int rank = a.ctl.rank;
  uint32_t* state = push_state_of(a.ctl.flags[rank]);
  const uint32_t step = state[0], prev_kind = state[3];
  const int64_t n_prev = state[2];
  const int cur = step % 3, prev = (step + 2) % 3;
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * kPeerThreads + threadIdx.x;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kPeerThreads;
  const int64_t cap = slot_vec / 2, part = n_vec / WORLD;
  const int64_t my_len = slice_len(n_vec, WORLD, rank), longest = slice_len(n_vec, WORLD, WORLD - 1);
  const int64_t buf_off = static_cast<int64_t>(cur) * WORLD * slot_vec;
  const uint4* in = reinterpret_cast<const uint4*>(a.in[rank]);
  uint4* out = reinterpret_cast<uint4*>(a.out);
  uint4* mine = reinterpret_cast<uint4*>(a.push[rank]);
  // 1. scatter: my copy of slice p goes to A[rank] of rank p
  for (int64_t j = tid; j < longest; j += stride) {
#pragma unroll
    for (int d = 1; d < WORLD; ++d) {
      const int p = (rank + d) % WORLD;
      if (j < (p == WORLD - 1 ? longest : part))
        st_peer_v4(reinterpret_cast<uint4*>(a.push[p]) + buf_off + rank * cap + j, canon_vec(in[p * part + j]));
    }
  }
  push_rearm<WORLD>(mine, prev, prev_kind, n_prev, slot_vec, rank, tid, stride);
  // 2. reduce my slice in rank order, keep it and post it to B[rank] of every peer
  const uint64_t t0 = globaltimer_ns();
  for (int64_t j = tid; j < my_len; j += stride) {
    uint4 v[WORLD];
    v[rank] = in[rank * part + j];
#pragma unroll
    for (int r = 0; r < WORLD; ++r)
      if (r != rank) v[r] = poll_vec(a, mine + buf_off + r * cap + j, t0, r, step);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < WORLD; ++r) acc_bf16x8(acc, v[r]);
    const uint4 sum = canon_vec(pack_bf16x8(acc));  // a sum that starts from +0 is never -0: canon is a no-op kept for safety
    out[rank * part + j] = sum;
#pragma unroll
    for (int d = 1; d < WORLD; ++d) {
      const int p = (rank + d) % WORLD;
      st_peer_v4(reinterpret_cast<uint4*>(a.push[p]) + buf_off + WORLD * cap + rank * cap + j, sum);
    }
  }
  // 3. gather: the other slices arrive reduced in B[r]
  for (int64_t j = tid; j < longest; j += stride) {
#pragma unroll
    for (int d = 1; d < WORLD; ++d) {
      const int r = (rank + d) % WORLD;
      if (j < (r == WORLD - 1 ? longest : part))
        out[r * part + j] = poll_vec(a, mine + buf_off + WORLD * cap + r * cap + j, t0, r, step);
    }
  }
  push_advance(state, step, n_vec, 2u);
}

// NVLS: the switch reduces. mc_in / mc_out are MULTICAST addresses of the input / output regions;
// the result lands in every rank's output region (a.out must be this rank's unicast view of it).
__global__ void __launch_bounds__(kPeerThreads) peer_allreduce_multimem_kernel(PeerArgs a, int64_t n_vec) {
  pdl_prologue();
  uint32_t epoch = *epoch_of(a.ctl.flags[a.ctl.rank], blockIdx.x);
  const int64_t part = n_vec / a.ctl.world;
  const int64_t my_start = a.ctl.rank * part;
  const int64_t my_len = (a.ctl.rank == a.ctl.world - 1) ? n_vec - my_start : part;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kPeerThreads;
  peer_barrier(a.ctl, epoch, false);
  const uint4* src = reinterpret_cast<const uint4*>(a.mc_in) + my_start;
  uint4* dst = reinterpret_cast<uint4*>(a.mc_out) + my_start;
  for (int64_t j = static_cast<int64_t>(blockIdx.x) * kPeerThreads + threadIdx.x; j < my_len; j += stride) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(src + j)
                 : "memory");
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + j), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
  }
  peer_barrier(a.ctl, epoch, /*fence=*/true);             // every slice has been broadcast
  if (threadIdx.x == 0) *epoch_of(a.ctl.flags[a.ctl.rank], blockIdx.x) = epoch;
}

template <typename K, typename... Args>
static cudaError_t peer_launch(K kern, int grid, bool pdl, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(kPeerThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, args...);
}

template <int WORLD>
static cudaError_t launch_world(const PeerArgs& a, int64_t n_vec, int algo, int grid, bool pdl, cudaStream_t stream) {
  if (algo == kPeerOneShot) return peer_launch(peer_allreduce_oneshot_kernel<WORLD>, grid, pdl, stream, a, n_vec);
  return peer_launch(peer_allreduce_twoshot_kernel<WORLD>, grid, pdl, stream, a, n_vec);
}

template <int WORLD>
static cudaError_t launch_push(const PeerArgs& a, int64_t n_vec, int grid, bool pdl, cudaStream_t stream) {
  return peer_launch(peer_allreduce_push_kernel<WORLD>, grid, pdl, stream, a, n_vec, a.push_slot_vec);
}

cudaError_t launch_peer_allreduce(const PeerArgs& a, int64_t n_vec, int algo, int grid, bool pdl, cudaStream_t stream) {
  if (algo == kPeerPush2) {
    switch (a.ctl.world) {
      case 2: return peer_launch(peer_allreduce_push2_kernel<2>, grid, pdl, stream, a, n_vec, a.push_slot_vec);
      case 4: return peer_launch(peer_allreduce_push2_kernel<4>, grid, pdl, stream, a, n_vec, a.push_slot_vec);
      case 8: return peer_launch(peer_allreduce_push2_kernel<8>, grid, pdl, stream, a, n_vec, a.push_slot_vec);
      default: return cudaErrorInvalidValue;
    }
  }
  if (algo == kPeerPush) {
    switch (a.ctl.world) {
      case 1: return launch_push<1>(a, n_vec, grid, pdl, stream);
      case 2: return launch_push<2>(a, n_vec, grid, pdl, stream);
      case 4: return launch_push<4>(a, n_vec, grid, pdl, stream);
      case 8: return launch_push<8>(a, n_vec, grid, pdl, stream);
      default: return cudaErrorInvalidValue;
    }
  }
  if (algo == kPeerMultimem) return peer_launch(peer_allreduce_multimem_kernel, grid, pdl, stream, a, n_vec);
  switch (a.ctl.world) {
    case 1: return launch_world<1>(a, n_vec, algo, grid, pdl, stream);
    case 2: return launch_world<2>(a, n_vec, algo, grid, pdl, stream);
    case 4: return launch_world<4>(a, n_vec, algo, grid, pdl, stream);
    case 8: return launch_world<8>(a, n_vec, algo, grid, pdl, stream);
    default: return cudaErrorInvalidValue;
  }
}

// ---------------------------------------------------------------------------------------------
// Row exchange inside a KV-head replica group (see PeerExchangeArgs in kernels.h): the q rows are
// PULLED from the peers after one flag barrier (their producers have finished), the output rows are
// PUSHED into the peers' buffers and published by one fenced barrier. One 16-byte vector per thread
// and step; vector v of the local staging [nrows][R][row_vec] maps to (i, g, c).
// ---------------------------------------------------------------------------------------------
template <bool SCATTER>
__global__ void __launch_bounds__(kPeerThreads) peer_exchange_kernel(PeerExchangeArgs a) {
  pdl_prologue();
  uint32_t epoch = *epoch_of(a.ctl.flags[a.ctl.rank], blockIdx.x);
  if (!SCATTER) peer_barrier(a.ctl, epoch, /*fence=*/false);
  const int R = a.ctl.world;
  const int64_t total = static_cast<int64_t>(a.nrows) * R * a.row_vec;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kPeerThreads;
  uint4* local = reinterpret_cast<uint4*>(a.local_rows);
  for (int64_t v = static_cast<int64_t>(blockIdx.x) * kPeerThreads + threadIdx.x; v < total; v += stride) {
    const int c = static_cast<int>(v % a.row_vec);
    const int64_t ig = v / a.row_vec;
    const int g = static_cast<int>(ig % R);
    const int64_t i = ig / R;
    uint4* remote = reinterpret_cast<uint4*>(a.peer_rows[g]) + (a.b0 + i) * a.row_vec + c;
    if (SCATTER) {
      const uint4 x = local[v];
      asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(remote), "r"(x.x), "r"(x.y), "r"(x.z),
                   "r"(x.w)
                   : "memory");
    } else {
      local[v] = ld_peer_v4(remote);
    }
  }
  if (SCATTER) peer_barrier(a.ctl, epoch, /*fence=*/true);  // my rows have landed in every peer's buffer
  if (threadIdx.x == 0) *epoch_of(a.ctl.flags[a.ctl.rank], blockIdx.x) = epoch;
}

cudaError_t launch_peer_exchange(const PeerExchangeArgs& a, bool scatter, int grid, bool pdl, cudaStream_t stream) {
  if (scatter) return peer_launch(peer_exchange_kernel<true>, grid, pdl, stream, a);
  return peer_launch(peer_exchange_kernel<false>, grid, pdl, stream, a);
}

}  // namespace b200
