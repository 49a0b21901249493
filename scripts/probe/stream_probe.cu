// Stand-alone HBM read-stream probe for the paged KV layout (diagnostic, not part of the product).
// Streams `npages` random 8 KiB (K page + V page of one (block, kv head)) units per launch with
// different staging methods and reports GB/s. Used to pick the staging method of the decode kernel.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_probe stream_probe.cu -lcuda
//   ./stream_probe [total_MiB=2048]
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <numeric>
#include <random>
#include <vector>

#include "../../vllm_omni_b200/csrc/ptx.cuh"
using namespace b200;

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e_ = (x);                                                          \
    if (e_ != cudaSuccess) {                                                       \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

constexpr int kPage = 8192;  // bytes per (block, head): K 4 KiB + V 4 KiB

// bounded mbarrier wait: a probe must never hang the box
__device__ __forceinline__ void mbar_wait_dbg(uint64_t* bar, uint32_t parity, int tag, int s) {
  for (long long i = 0; i < 2000000; ++i)
    if (mbar_try_wait(bar, parity)) return;
  printf("mbar timeout tag=%d block=%d warp=%d s=%d parity=%u\n", tag, blockIdx.x, threadIdx.x >> 5, s, parity);
  __trap();
}

__device__ __forceinline__ void bulk_copy_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar,
                                             uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d_nohint(void* dst, const CUtensorMap* tm, uint64_t* bar, int c0,
                                                   int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// ---------------------------------------------------------------------------------------------
// mode 0: LDG.128 straight to registers. Each warp takes whole pages; U loads in flight / thread.
// ---------------------------------------------------------------------------------------------
template <int U>
__global__ void __launch_bounds__(256) k_ldg(const uint8_t* base, const int* pages, int npages, int* sink) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  int acc = 0;
  for (int p = warp; p < npages; p += nwarps) {
    const uint8_t* src = base + static_cast<int64_t>(pages[p]) * kPage;
#pragma unroll
    for (int i = 0; i < kPage / 512; i += U) {
      int4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) v[u] = ldg_nc_v4(src + (i + u) * 512 + lane * 16);
#pragma unroll
      for (int u = 0; u < U; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
  }
  if (acc == 0x12345678) *sink = acc;
}

// ---------------------------------------------------------------------------------------------
// mode 1: cp.async 16 B into a per-warp ring of D stages of SP pages; persistent CTAs.
// ---------------------------------------------------------------------------------------------
template <int W, int D, int SP>
__global__ void __launch_bounds__(W * 32) k_cpasync(const uint8_t* base, const int* pages, int npages,
                                                    int* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* ring = smem + warp * D * SP * kPage;
  const int gw = blockIdx.x * W + warp, nw = gridDim.x * W;
  const int nst = (npages / SP - gw + nw - 1) / nw;  // stages of this warp
  auto issue = [&](int s) {
    if (s < nst) {
      uint8_t* dst = ring + (s % D) * SP * kPage;
#pragma unroll
      for (int pg = 0; pg < SP; ++pg) {
        const uint8_t* src = base + static_cast<int64_t>(pages[(gw + s * nw) * SP + pg]) * kPage;
#pragma unroll
        for (int i = 0; i < kPage / 512; ++i) cp_async_16(dst + pg * kPage + i * 512 + lane * 16, src + i * 512 + lane * 16);
      }
    }
    cp_async_commit();
  };
  for (int s = 0; s < D - 1; ++s) issue(s);
  int acc = 0;
  for (int s = 0; s < nst; ++s) {
    issue(s + D - 1);
    cp_async_wait<D - 1>();
    __syncwarp();
    acc ^= *reinterpret_cast<const int*>(ring + (s % D) * SP * kPage + lane * 16);
    __syncwarp();
  }
  if (acc == 0x12345678) *sink = acc;
}

// ---------------------------------------------------------------------------------------------
// modes 2-4: one producer thread fills an mbarrier ring of NS stages (SP pages each); W consumer
// warps take stages round-robin, touch the data and release the slot.
//   2: cp.async.bulk 1-D, one 8 KiB op per page
//   3: TMA 4-D tile, 64x16 boxes (2 KiB per op, 4 ops per page)      [what the decode kernel does]
//   4: TMA 4-D tile, 64x32x2 box over the merged (kv,slot) rows (8 KiB per op)
// ---------------------------------------------------------------------------------------------
template <int MODE, int W, int NS, int SP>
__global__ void __launch_bounds__((W + 1) * 32, 1)
k_ring(const __grid_constant__ CUtensorMap tm, const uint8_t* base, const int* pages, int npages, int* sink) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + NS * SP * kPage);
  uint64_t* empty = full + NS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    fence_barrier_init();
  }
  __syncthreads();
  const int nst = (npages / SP - blockIdx.x + gridDim.x - 1) / gridDim.x;
  if (warp == 0) {
    const uint64_t pol = l2_policy_evict_first();
    // page ids are prefetched 32 stages at a time (lane L holds the ids of stage s0 + L), double
    // buffered, so the producer never waits on a dependent global load
    int ids[SP], ids_next[SP];
    auto fetch = [&](int s0, int (&dst)[SP]) {
#pragma unroll
      for (int j = 0; j < SP; ++j) {
        const int sj = s0 + lane;
        dst[j] = sj < nst ? pages[(blockIdx.x + sj * gridDim.x) * SP + j] : 0;
      }
    };
    fetch(0, ids);
    fetch(32, ids_next);
    for (int s = 0; s < nst; ++s) {
      const int slot = s % NS;
      if ((s & 31) == 0 && s > 0) {
#pragma unroll
        for (int j = 0; j < SP; ++j) ids[j] = ids_next[j];
        fetch(s + 32, ids_next);
      }
      int pgs[SP];
#pragma unroll
      for (int j = 0; j < SP; ++j) pgs[j] = __shfl_sync(0xffffffffu, ids[j], s & 31);
      if (lane == 0) {
        mbar_wait_dbg(&empty[slot], ((s / NS) & 1) ^ 1, 0, s);
        mbar_arrive_expect_tx(&full[slot], SP * kPage);
      }
      __syncwarp();
      uint8_t* dst = smem + slot * SP * kPage;
      if (MODE == 2) {
        if (lane < SP) {
          int pg = pgs[0];
#pragma unroll
          for (int j = 1; j < SP; ++j) pg = lane == j ? pgs[j] : pg;
          bulk_copy_1d(dst + lane * kPage, base + static_cast<int64_t>(pg) * kPage, kPage, &full[slot], pol);
        }
      } else if (MODE == 3) {
        if (lane < SP * 4) {
          int pg = pgs[0];
#pragma unroll
          for (int j = 1; j < SP; ++j) pg = (lane >> 2) == j ? pgs[j] : pg;
          const int kv = (lane >> 1) & 1, half = lane & 1;
          tma_load_4d(dst + lane * 2048, &tm, &full[slot], half * 64, kv * 16, 0, pg, pol);
        }
      } else {
        if (lane < SP) {
          int pg = pgs[0];
#pragma unroll
          for (int j = 1; j < SP; ++j) pg = lane == j ? pgs[j] : pg;
          tma_load_4d(dst + lane * kPage, &tm, &full[slot], 0, 0, 0, pg, pol);
        }
      }
    }
  } else {
    const int cw = warp - 1;
    int acc = 0;
    for (int s = cw; s < nst; s += W) {
      const int slot = s % NS;
      mbar_wait_dbg(&full[slot], (s / NS) & 1, 1, s);
      acc ^= *reinterpret_cast<const int*>(smem + slot * SP * kPage + lane * 16);
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
    }
    if (acc == 0x12345678) *sink = acc;
  }
}

// ---------------------------------------------------------------------------------------------
// modes 5/6: every warp fills its OWN ring (D slots of one 8 KiB page, one mbarrier per slot) with
// one bulk op per page issued by its lane 0, waits for the oldest, touches it and re-issues: no
// producer warp, no empty barriers, no cross-warp handshake.
//   5: cp.async.bulk 1-D (linear 8 KiB)      6: TMA 4-D tile 64x32x2 box (SWIZZLE_128B)
// ---------------------------------------------------------------------------------------------
template <int MODE, int W, int D, int DELAY = 0>
__global__ void __launch_bounds__(W * 32, 1)
k_self(const __grid_constant__ CUtensorMap tm, const uint8_t* base, const int* pages, int npages, int* sink) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + W * D * kPage);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0)
    for (int s = 0; s < D; ++s) mbar_init(&bars[warp * D + s], 1);
  fence_barrier_init();
  __syncthreads();
  uint8_t* ring = smem + warp * D * kPage;
  uint64_t* bar = bars + warp * D;
  const int gw = blockIdx.x * W + warp, nw = gridDim.x * W;
  const int nst = (npages - gw + nw - 1) / nw;
  const uint64_t pol = l2_policy_evict_first();
  // page ids of the next 32 stages, one per lane
  int ids = (lane < nst) ? pages[gw + lane * nw] : 0;
  auto issue = [&](int s) {
    if (s < nst) {
      if ((s & 31) == 0 && s > 0) ids = (s + lane < nst) ? pages[gw + (s + lane) * nw] : 0;
      const int pg = __shfl_sync(0xffffffffu, ids, s & 31);
      if (lane == 0) {
        const int slot = s % D;
        mbar_arrive_expect_tx(&bar[slot], kPage);
        if (MODE == 5)
          bulk_copy_1d(ring + slot * kPage, base + static_cast<int64_t>(pg) * kPage, kPage, &bar[slot], pol);
        else if (MODE == 6)
          tma_load_4d(ring + slot * kPage, &tm, &bar[slot], 0, 0, 0, pg, pol);
        else {  // 7: two 4 KiB boxes (K rows 0-15, V rows 16-31), as the decode kernel issues them
          tma_load_4d(ring + slot * kPage, &tm, &bar[slot], 0, 0, 0, pg, pol);
          tma_load_4d(ring + slot * kPage + 4096, &tm, &bar[slot], 0, 16, 0, pg, pol);
        }
      }
    }
  };
  for (int s = 0; s < D; ++s) issue(s);
  int acc = 0;
  for (int s = 0; s < nst; ++s) {
    const int slot = s % D;
    mbar_wait_dbg(&bar[slot], (s / D) & 1, 2, s);
    acc ^= *reinterpret_cast<const int*>(ring + slot * kPage + lane * 16);
    if (DELAY > 0) {  // stand-in for the per-tile math of the decode kernel
      const long long t0 = clock64();
      while (clock64() - t0 < DELAY) {
      }
    }
    __syncwarp();
    issue(s + D);
  }
  if (acc == 0x12345678) *sink = acc;
}

// ---------------------------------------------------------------------------------------------
// mode 8: as mode 7 (self-issued 2 x 4 KiB TMA boxes per stage) but the warp consumes whichever
// slot has landed first (softmax accumulation is order independent inside a work item), so one
// slow DRAM request does not block the stages behind it.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

template <int W, int D, int DELAY>
__global__ void __launch_bounds__(W * 32, 1)
k_ooo(const __grid_constant__ CUtensorMap tm, const uint8_t* base, const int* pages, int npages, int* sink) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + W * D * kPage);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0)
    for (int s = 0; s < D; ++s) mbar_init(&bars[warp * D + s], 1);
  fence_barrier_init();
  __syncthreads();
  uint8_t* ring = smem + warp * D * kPage;
  uint64_t* bar = bars + warp * D;
  const int gw = blockIdx.x * W + warp, nw = gridDim.x * W;
  const int nst = (npages - gw + nw - 1) / nw;
  const uint64_t pol = l2_policy_evict_first();
  int ids = (lane < nst) ? pages[gw + lane * nw] : 0;
  int next = 0;            // next stage to issue
  uint32_t phases = 0;     // bit s = parity slot s waits for
  uint32_t busy = 0;       // bit s = slot s holds an issued, unconsumed stage
  auto issue = [&](int slot) {
    if (next < nst) {
      if ((next & 31) == 0 && next > 0) ids = (next + lane < nst) ? pages[gw + (next + lane) * nw] : 0;
      const int pg = __shfl_sync(0xffffffffu, ids, next & 31);
      if (lane == 0) {
        mbar_arrive_expect_tx(&bar[slot], kPage);
        tma_load_4d(ring + slot * kPage, &tm, &bar[slot], 0, 0, 0, pg, pol);
        tma_load_4d(ring + slot * kPage + 4096, &tm, &bar[slot], 0, 16, 0, pg, pol);
      }
      busy |= 1u << slot;
      ++next;
    }
  };
  for (int s = 0; s < D; ++s) issue(s);
  int acc = 0, oldest = 0;
  for (int done = 0; done < nst; ++done) {
    int slot = -1;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      const int s = (oldest + i) % D;
      if (slot < 0 && ((busy >> s) & 1) && mbar_test(&bar[s], (phases >> s) & 1)) slot = s;
    }
    slot = __shfl_sync(0xffffffffu, slot, 0);
    if (slot < 0) {  // nothing landed yet: block on the oldest busy slot
      slot = oldest;
      while (!((busy >> slot) & 1)) slot = (slot + 1) % D;
      mbar_wait_dbg(&bar[slot], (phases >> slot) & 1, 3, done);
    } else {
      mbar_wait_dbg(&bar[slot], (phases >> slot) & 1, 4, done);  // make it visible to every lane
    }
    acc ^= *reinterpret_cast<const int*>(ring + slot * kPage + lane * 16);
    if (DELAY > 0) {
      const long long t0 = clock64();
      while (clock64() - t0 < DELAY) {
      }
    }
    __syncwarp();
    phases ^= 1u << slot;
    busy &= ~(1u << slot);
    if (slot == oldest) oldest = (oldest + 1) % D;
    issue(slot);
  }
  if (acc == 0x12345678) *sink = acc;
}

// ---------------------------------------------------------------------------------------------
// mode 9: as mode 7, plus an L2 prefetch (cp.async.bulk.prefetch.tensor, no shared memory, no
// completion) issued PD stages ahead of the shared-memory load: DRAM latency is covered by the
// L2-resident window, the ring only has to cover the L2 -> SM hop.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_4d(const CUtensorMap* tm, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.prefetch.tensor.4d.L2.global.tile [%0, {%1, %2, %3, %4}];" ::"l"(
                   reinterpret_cast<uint64_t>(tm)),
               "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

template <int W, int D, int PD, int DELAY>
__global__ void __launch_bounds__(W * 32, 1)
k_l2pf(const __grid_constant__ CUtensorMap tm, const __grid_constant__ CUtensorMap tm_pf, const uint8_t* base,
       const int* pages, int npages, int* sink) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + W * D * kPage);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0)
    for (int s = 0; s < D; ++s) mbar_init(&bars[warp * D + s], 1);
  fence_barrier_init();
  __syncthreads();
  uint8_t* ring = smem + warp * D * kPage;
  uint64_t* bar = bars + warp * D;
  const int gw = blockIdx.x * W + warp, nw = gridDim.x * W;
  const int nst = (npages - gw + nw - 1) / nw;
  const uint64_t pol = l2_policy_evict_first();
  auto page_of = [&](int s) { return pages[gw + s * nw]; };  // (probe only: plain dependent load)
  auto prefetch = [&](int s) {
    if (s < nst && lane == 0) tma_prefetch_4d(&tm_pf, 0, 0, 0, page_of(s));
  };
  auto issue = [&](int s) {
    if (s < nst && lane == 0) {
      const int slot = s % D;
      const int pg = page_of(s);
      mbar_arrive_expect_tx(&bar[slot], kPage);
      tma_load_4d(ring + slot * kPage, &tm, &bar[slot], 0, 0, 0, pg, pol);
      tma_load_4d(ring + slot * kPage + 4096, &tm, &bar[slot], 0, 16, 0, pg, pol);
    }
  };
  for (int s = 0; s < PD; ++s) prefetch(s);
  for (int s = 0; s < D; ++s) issue(s);
  int acc = 0;
  for (int s = 0; s < nst; ++s) {
    const int slot = s % D;
    prefetch(s + PD);
    mbar_wait_dbg(&bar[slot], (s / D) & 1, 5, s);
    acc ^= *reinterpret_cast<const int*>(ring + slot * kPage + lane * 16);
    if (DELAY > 0) {
      const long long t0 = clock64();
      while (clock64() - t0 < DELAY) {
      }
    }
    __syncwarp();
    issue(s + D);
  }
  if (acc == 0x12345678) *sink = acc;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_enc() {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &p, 12000, cudaEnableDefault, &q));
  return reinterpret_cast<EncodeTiledFn>(p);
}

template <typename F>
static void run(const char* name, size_t bytes, F&& launch) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int i = 0; i < 3; ++i) launch();
  CK(cudaDeviceSynchronize());
  const int reps = 10;
  CK(cudaEventRecord(e0));
  for (int i = 0; i < reps; ++i) launch();
  CK(cudaEventRecord(e1));
  cudaError_t es = cudaDeviceSynchronize();
  if (es != cudaSuccess) {
    printf("%-44s FAILED: %s\n", name, cudaGetErrorString(es));
    exit(2);
  }
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  printf("%-44s %8.1f us  %8.1f GB/s\n", name, ms / reps * 1e3, bytes / (ms / reps * 1e-3) / 1e9);
  fflush(stdout);
}

int main(int argc, char** argv) {
  const size_t total_mib = argc > 1 ? atoi(argv[1]) : 2048;
  const size_t pool_pages = total_mib * 1024 * 1024 / kPage * 4;  // pool 4x larger than one pass
  const int npages = static_cast<int>(total_mib * 1024 * 1024 / kPage);
  uint8_t* base;
  CK(cudaMalloc(&base, pool_pages * kPage));
  CK(cudaMemset(base, 1, pool_pages * kPage));
  std::vector<int> perm(pool_pages);
  std::iota(perm.begin(), perm.end(), 0);
  std::mt19937 rng(0);
  std::shuffle(perm.begin(), perm.end(), rng);
  int* pages;
  CK(cudaMalloc(&pages, pool_pages * sizeof(int)));
  CK(cudaMemcpy(pages, perm.data(), pool_pages * sizeof(int), cudaMemcpyHostToDevice));
  int* sink;
  CK(cudaMalloc(&sink, 4));
  int sms;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const size_t bytes = static_cast<size_t>(npages) * kPage;
  printf("SMs %d, %d pages of %d B per launch (%.0f MiB), pool %.0f MiB, random page order\n", sms, npages, kPage,
         bytes / 1048576.0, pool_pages * kPage / 1048576.0);
  int rot = 0;
  auto pg = [&]() { rot = (rot + 1) % 4; return pages + static_cast<size_t>(rot) * npages; };

  // tensor maps over the pool viewed as [page][32 rows (K 0-15, V 0-15)][128 bf16]
  EncodeTiledFn enc = get_enc();
  CUtensorMap tm3, tm4;
  {
    cuuint64_t gdim[4] = {128, 32, 1, pool_pages};
    cuuint64_t gstr[3] = {256, 8192, 8192};
    cuuint32_t box[4] = {64, 16, 1, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = enc(&tm3, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, base, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) printf("encode tm3 failed %d\n", r);
  }
  {
    // [64 (d_lo)][32 rows, stride 256 B][2 (d_hi), stride 128 B][page]
    cuuint64_t gdim[4] = {64, 32, 2, pool_pages};
    cuuint64_t gstr[3] = {256, 128, 8192};
    cuuint32_t box[4] = {64, 32, 2, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = enc(&tm4, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, base, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) printf("encode tm4 failed %d\n", r);
  }

  CUtensorMap tm7;
  {
    cuuint64_t gdim[4] = {64, 32, 2, pool_pages};
    cuuint64_t gstr[3] = {256, 128, 8192};
    cuuint32_t box[4] = {64, 16, 2, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = enc(&tm7, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, base, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) printf("encode tm7 failed %d\n", r);
  }
  run("ldg128 U=4, 8 CTA/SM x 256thr", bytes, [&] { k_ldg<4><<<sms * 8, 256>>>(base, pg(), npages, sink); });
  run("ldg128 U=8, 8 CTA/SM x 256thr", bytes, [&] { k_ldg<8><<<sms * 8, 256>>>(base, pg(), npages, sink); });
  run("ldg128 U=16, 4 CTA/SM x 256thr", bytes, [&] { k_ldg<16><<<sms * 4, 256>>>(base, pg(), npages, sink); });

#define CP(W, D, SP)                                                                              \
  {                                                                                               \
    auto k = k_cpasync<W, D, SP>;                                                                 \
    const int sm = W * D * SP * kPage;                                                            \
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));                 \
    run("cp.async16 W=" #W " depth=" #D " stage=" #SP "pg", bytes, [&] { k<<<sms, W * 32, sm>>>(base, pg(), npages, sink); }); \
  }
  CP(4, 2, 2);
  CP(4, 3, 2);
  CP(8, 2, 1);
  CP(8, 3, 1);
  CP(4, 6, 1);
  CP(12, 2, 1);

#define RING(MODE, W, NS, SP, TM)                                                                 \
  {                                                                                               \
    auto k = k_ring<MODE, W, NS, SP>;                                                             \
    const int sm = NS * SP * kPage + 2 * NS * 8 + 1024;                                           \
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));                 \
    run("ring mode=" #MODE " W=" #W " NS=" #NS " SP=" #SP, bytes, [&] { k<<<sms, (W + 1) * 32, sm>>>(TM, base, pg(), npages, sink); }); \
  }
  // NS must be a multiple of W: a consumer warp may only wait on a slot whose previous fill it
  // consumed itself, otherwise the 1-bit phase parity can alias when loads land out of order.
  RING(2, 8, 8, 2, tm3);
  RING(2, 8, 16, 1, tm3);
  RING(2, 8, 24, 1, tm3);
  RING(2, 4, 12, 2, tm3);
  RING(2, 4, 4, 4, tm3);
  RING(2, 2, 6, 4, tm3);
  RING(3, 8, 8, 2, tm3);
  RING(3, 8, 24, 1, tm3);
  RING(4, 8, 8, 2, tm4);
  RING(4, 8, 24, 1, tm4);
  RING(4, 4, 4, 4, tm4);
#define SELF(MODE, W, D, TM) SELFD(MODE, W, D, 0, TM)
#define SELFD(MODE, W, D, DELAY, TM)                                                              \
  {                                                                                               \
    auto k = k_self<MODE, W, D, DELAY>;                                                                  \
    const int sm = W * D * kPage + W * D * 8 + 1024;                                              \
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));                 \
    run("self mode=" #MODE " W=" #W " D=" #D " delay=" #DELAY, bytes, [&] { k<<<sms, W * 32, sm>>>(TM, base, pg(), npages, sink); }); \
  }
  SELF(5, 8, 2, tm4);
  SELF(5, 8, 3, tm4);
  SELF(5, 4, 6, tm4);
  SELF(5, 12, 2, tm4);
  SELF(6, 8, 2, tm4);
  SELF(6, 8, 3, tm4);
  SELF(6, 4, 6, tm4);
  SELF(6, 12, 2, tm4);
  SELFD(7, 8, 3, 0, tm7);
  SELFD(7, 8, 3, 500, tm7);
  SELFD(7, 8, 3, 1000, tm7);
  SELFD(7, 8, 3, 1500, tm7);
  SELFD(7, 8, 3, 2000, tm7);
  SELFD(7, 9, 3, 1500, tm7);
  SELFD(7, 6, 4, 1500, tm7);
  SELFD(7, 12, 2, 1500, tm7);
  SELFD(7, 13, 2, 1500, tm7);
  SELFD(7, 4, 6, 1500, tm7);
  SELFD(7, 5, 5, 1500, tm7);
#define OOO(W, D, DELAY)                                                                          \
  {                                                                                               \
    auto k = k_ooo<W, D, DELAY>;                                                                  \
    const int sm = W * D * kPage + W * D * 8 + 1024;                                              \
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));                 \
    run("ooo W=" #W " D=" #D " delay=" #DELAY, bytes, [&] { k<<<sms, W * 32, sm>>>(tm7, base, pg(), npages, sink); }); \
  }
#define L2PF(W, D, PD, DELAY)                                                                     \
  {                                                                                               \
    auto k = k_l2pf<W, D, PD, DELAY>;                                                             \
    const int sm = W * D * kPage + W * D * 8 + 1024;                                              \
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));                 \
    run("l2pf W=" #W " D=" #D " PD=" #PD " delay=" #DELAY, bytes, [&] { k<<<sms, W * 32, sm>>>(tm7, tm4, base, pg(), npages, sink); }); \
  }
  L2PF(8, 3, 6, 0);
  L2PF(8, 3, 6, 1500);
  L2PF(8, 3, 10, 1500);
  L2PF(8, 3, 16, 1500);
  L2PF(8, 2, 8, 1500);
  L2PF(12, 2, 8, 1500);
  L2PF(12, 2, 12, 1500);
  OOO(8, 3, 0);
  OOO(8, 3, 1000);
  OOO(8, 3, 1500);
  OOO(6, 4, 1500);
  OOO(4, 6, 1500);
  OOO(12, 2, 1500);
  return 0;
}
