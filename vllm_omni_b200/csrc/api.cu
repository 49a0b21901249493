// C ABI of libb200attn.so (include/b200attn.h): host-side validation, split-K heuristic,
// TMA tensor-map encoding and kernel launches. No torch types, no allocation, no sync.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/b200attn.h"
#include "kernels.h"

using namespace b200;

namespace {

thread_local char g_err[512] = "";

constexpr bool kPrefillUmmaDefault = true;  // parity-green on B200 (tests/test_gpu_prefill.py runs both kernels)

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return -code;
}

int cuda_fail(cudaError_t e, const char* what) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
  return static_cast<int>(e);
}

int cached_num_sms() {
  static thread_local int cached_dev = -1;
  static thread_local int cached_sms = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev != cached_dev) {
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    cached_dev = dev;
    cached_sms = sms;
  }
  return cached_sms;
}

bool env_flag(const char* name, bool dflt) {
  const char* v = getenv(name);
  if (v == nullptr || *v == '\0') return dflt;
  return !(v[0] == '0' || v[0] == 'n' || v[0] == 'N' || v[0] == 'f' || v[0] == 'F');
}

// ---------------------------------------------------------------------------------------------
// tensor maps
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &p, 12000, cudaEnableDefault,
                                         &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

struct KvMapKey {
  const void* ptr;
  int64_t sb, ss, sh, nb;
  int32_t hkv, bs, fp8;
  bool operator==(const KvMapKey& o) const {
    return ptr == o.ptr && sb == o.sb && ss == o.ss && sh == o.sh && nb == o.nb && hkv == o.hkv &&
           bs == o.bs && fp8 == o.fp8;
  }
};
struct KvMapEntry {
  unsigned long long stamp;  // last use (least recently used way of a set is replaced)
  KvMapKey key;
  CUtensorMap map;
  CoordPerm perm;
  bool valid;
};
constexpr int kMapCache = 1024;  // >> 2 maps x layers of a model; 8-way set-associative by pointer hash, LRU within a set
constexpr int kMapWays = 8;
thread_local KvMapEntry g_maps[kMapCache];
thread_local unsigned long long g_map_encodes = 0;  // tensor maps encoded by this thread (a miss costs ~1 us of host time)
thread_local unsigned long long g_map_clock = 0;

// The cached map for `key`, or the entry to (re)encode into: a free way of the key's set, else its least recently used.
KvMapEntry* find_map(KvMapEntry* table, const KvMapKey& key, bool* hit) {
  uint64_t h = reinterpret_cast<uintptr_t>(key.ptr) >> 8;
  h = (h ^ (h >> 17)) * 0x9E3779B97F4A7C15ull;
  const size_t set = (h >> 40) % (kMapCache / kMapWays) * kMapWays;
  KvMapEntry* victim = nullptr;
  for (int w = 0; w < kMapWays; ++w) {
    KvMapEntry& e = table[set + w];
    if (e.valid && e.key == key) {
      *hit = true;
      e.stamp = ++g_map_clock;
      return &e;
    }
    if (victim == nullptr || (!e.valid && victim->valid) || (e.valid == victim->valid && e.stamp < victim->stamp)) victim = &e;
  }
  *hit = false;
  ++g_map_encodes;
  victim->stamp = ++g_map_clock;
  return victim;
}

// Encode (or fetch) the 4-D tensor map of one cache tensor: dims sorted by stride
// (head_size, then slot/head/block in ascending stride order), box = 64 x 16 tokens x 1 x 1
// (128 x 16 for e4m3), SWIZZLE_128B.
int get_kv_map(const void* ptr, int64_t sb, int64_t ss, int64_t sh, int64_t nb, int32_t hkv,
               int32_t bs, int32_t hd, bool fp8, const CUtensorMap** out_map, CoordPerm* out_perm) {
  KvMapKey key{ptr, sb, ss, sh, nb, hkv, bs, fp8 ? 1 : 0};
  bool hit = false;
  KvMapEntry& e = *find_map(g_maps, key, &hit);
  if (hit) {
    *out_map = &e.map;
    *out_perm = e.perm;
    return 0;
  }
  e.valid = false;
  EncodeTiledFn enc = get_encode_fn();
  if (enc == nullptr) return fail(B200ATTN_EDRIVER, "cuTensorMapEncodeTiled entry point not found");
  const int64_t es = fp8 ? 1 : 2;
  // logical dims: 0 = slot, 1 = head, 2 = block
  int64_t size[3] = {bs, hkv, nb};
  int64_t stride[3] = {ss * es, sh * es, sb * es};
  // a size-1 dimension may carry any stride: give it one that keeps the order canonical
  if (hkv == 1) stride[1] = stride[0] * bs;
  if (nb == 1) stride[2] = (stride[1] * hkv > stride[0] * bs) ? stride[1] * hkv : stride[0] * bs;
  int order[3] = {0, 1, 2};
  for (int i = 0; i < 3; ++i)
    for (int j = i + 1; j < 3; ++j)
      if (stride[order[j]] < stride[order[i]]) {
        int t = order[i];
        order[i] = order[j];
        order[j] = t;
      }
  cuuint64_t gdim[4] = {static_cast<cuuint64_t>(hd), 0, 0, 0};
  cuuint64_t gstr[3] = {0, 0, 0};
  cuuint32_t box[4] = {static_cast<cuuint32_t>(fp8 ? 128 : 64), 1, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  for (int i = 0; i < 3; ++i) {
    const int d = order[i];
    gdim[i + 1] = static_cast<cuuint64_t>(size[d]);
    gstr[i] = static_cast<cuuint64_t>(stride[d]);
    if (stride[d] <= 0 || (stride[d] % 16) != 0)
      return fail(B200ATTN_EINVAL, "cache stride %lld B of dim %d is not a positive multiple of 16",
                  static_cast<long long>(stride[d]), d);
    if (d == 0) box[i + 1] = kTileTokens;
  }
  CUresult r = enc(&e.map, fp8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                   const_cast<void*>(ptr), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    e.valid = false;
    return fail(B200ATTN_EDRIVER,
                "cuTensorMapEncodeTiled failed (%d): dims [%llu,%llu,%llu,%llu] strides [%llu,%llu,%llu]",
                static_cast<int>(r), (unsigned long long)gdim[0], (unsigned long long)gdim[1],
                (unsigned long long)gdim[2], (unsigned long long)gdim[3], (unsigned long long)gstr[0],
                (unsigned long long)gstr[1], (unsigned long long)gstr[2]);
  }
  e.key = key;
  e.perm = CoordPerm{order[0], order[1], order[2]};
  e.valid = true;
  *out_map = &e.map;
  *out_perm = e.perm;
  return 0;
}

// 5-D tensor map of one cache tensor for the stream-K kernel, valid for ANY cache strides:
//   dim0 = d_lo (64 elements = 128 B, the swizzle span)   dim1 = slot   (stride cs_slot)
//   dim2 = d_hi (2, stride 64 elements)                   dim3 = head   (stride cs_head)
//   dim4 = block (stride cs_block)
// box = 64 x 16 x 2 x 1 x 1: one 16-token x 128-dim tile per op, landing as two 2 KiB halves of
// SWIZZLE_128B rows — the layout the math reads with ldmatrix.
thread_local KvMapEntry g_maps5[kMapCache];
thread_local KvMapEntry g_maps5h[kMapCache];  // bf16, box of ONE 64-element half (UMMA operand staging)

// box_halves = 2: one op moves both 64-element halves of a 16-token tile (4 KiB, the layout ldmatrix reads);
// box_halves = 1: one op moves one half (2 KiB), so a kernel can place the halves of many tiles into two
// [tokens][128 B] regions — the SWIZZLE_128B operand layout tcgen05.mma descriptors describe.
int get_kv_map5(const void* ptr, int64_t sb, int64_t ss, int64_t sh, int64_t nb, int32_t hkv, int32_t bs, bool fp8,
                const CUtensorMap** out_map, int box_halves = 2) {
  KvMapKey key{ptr, sb, ss, sh, nb, hkv, bs, fp8 ? 1 : (box_halves == 1 ? 2 : 0)};
  bool hit = false;
  KvMapEntry& e = *find_map(box_halves == 1 ? g_maps5h : g_maps5, key, &hit);
  if (hit) {
    *out_map = &e.map;
    return 0;
  }
  e.valid = false;
  EncodeTiledFn enc = get_encode_fn();
  if (enc == nullptr) return fail(B200ATTN_EDRIVER, "cuTensorMapEncodeTiled entry point not found");
  if (fp8) {
    // e4m3 cache: a 128-element row is exactly the 128 B swizzle span -> 4-D map
    // [128][slot][head][block], box 128 x 16: one 2 KiB tile per op
    const int64_t f_head = hkv > 1 ? sh : 128, f_block = nb > 1 ? sb : 128;
    cuuint64_t gdim[4] = {128, static_cast<cuuint64_t>(bs), static_cast<cuuint64_t>(hkv), static_cast<cuuint64_t>(nb)};
    cuuint64_t gstr[3] = {static_cast<cuuint64_t>(ss), static_cast<cuuint64_t>(f_head), static_cast<cuuint64_t>(f_block)};
    cuuint32_t box[4] = {128, static_cast<cuuint32_t>(kTileTokens), 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&e.map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(ptr), gdim, gstr, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      e.valid = false;
      return fail(B200ATTN_EDRIVER, "cuTensorMapEncodeTiled (fp8 4-D) failed (%d): strides [%lld,%lld,%lld] B",
                  static_cast<int>(r), static_cast<long long>(ss), static_cast<long long>(f_head),
                  static_cast<long long>(f_block));
    }
    e.key = key;
    e.valid = true;
    *out_map = &e.map;
    return 0;
  }
  // a size-1 dimension may carry any stride: give it a harmless one
  const int64_t st_slot = ss * 2;
  const int64_t st_head = hkv > 1 ? sh * 2 : 256;
  const int64_t st_block = nb > 1 ? sb * 2 : 256;
  cuuint64_t gdim[5] = {64, static_cast<cuuint64_t>(bs), 2, static_cast<cuuint64_t>(hkv), static_cast<cuuint64_t>(nb)};
  cuuint64_t gstr[4] = {static_cast<cuuint64_t>(st_slot), 128, static_cast<cuuint64_t>(st_head),
                        static_cast<cuuint64_t>(st_block)};
  cuuint32_t box[5] = {64, static_cast<cuuint32_t>(kTileTokens), static_cast<cuuint32_t>(box_halves), 1, 1};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(&e.map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    e.valid = false;
    return fail(B200ATTN_EDRIVER, "cuTensorMapEncodeTiled (5-D) failed (%d): strides [%lld,128,%lld,%lld] B",
                static_cast<int>(r), static_cast<long long>(st_slot), static_cast<long long>(st_head),
                static_cast<long long>(st_block));
  }
  e.key = key;
  e.valid = true;
  *out_map = &e.map;
  return 0;
}

int check_cache_args(const void* k_cache, const void* v_cache, int64_t sb, int64_t ss, int64_t sh,
                     int64_t nb, int32_t hkv, int32_t hd, int32_t bs, int32_t kv_dtype) {
  if (k_cache == nullptr || v_cache == nullptr) return fail(B200ATTN_EINVAL, "null cache pointer");
  if (kv_dtype != B200ATTN_KV_BF16 && kv_dtype != B200ATTN_KV_FP8_E4M3)
    return fail(B200ATTN_EINVAL, "kv_dtype %d unknown", kv_dtype);
  if (hd != kHeadDim) return fail(B200ATTN_EUNSUPPORTED, "head_size %d not built (128 only)", hd);
  if (bs <= 0 || (bs % 16) != 0)
    return fail(B200ATTN_EUNSUPPORTED, "block_size %d is not a multiple of 16", bs);
  if (hkv <= 0 || nb <= 0) return fail(B200ATTN_EINVAL, "num_kv_heads/num_blocks must be positive");
  const int64_t es = kv_dtype == B200ATTN_KV_FP8_E4M3 ? 1 : 2;
  if (ss <= 0 || (hkv > 1 && sh <= 0) || (nb > 1 && sb <= 0))
    return fail(B200ATTN_EINVAL, "cache strides must be positive");
  if (((ss * es) % 16) || (hkv > 1 && ((sh * es) % 16)) || (nb > 1 && ((sb * es) % 16)))
    return fail(B200ATTN_EINVAL, "cache strides must be multiples of 16 bytes");
  if ((reinterpret_cast<uintptr_t>(k_cache) % 16) || (reinterpret_cast<uintptr_t>(v_cache) % 16))
    return fail(B200ATTN_EINVAL, "cache pointers must be 16-byte aligned");
  return 0;
}

// stream-K workspace: 2 partial slots per warp of the persistent grid (O rows then LSE), then the
// tiles-prefix plan.
// Control area at the head of every decode workspace, zeroed once by b200attn_workspace_init; counters in
// it are left at zero by every kernel: [0, 256) stream-K v1 words, [256, ...) Sk2Ctrl of stream-K v2
// (arrival counters, chunk-group claims, per-SM rates and weights). No variant writes anything else here,
// so one workspace can serve different variants in turn.
constexpr size_t kWsCtrlBytes = 65536;
constexpr size_t kWsSk3CntOff = 16384;  // cooperative stream-K arrival counters [2 parities][teams <= 512 * 12]
constexpr size_t kWsSk2CtrlOff = 256;
static_assert(kWsSk2CtrlOff + sizeof(Sk2Ctrl) <= kWsSk3CntOff, "control area too small for Sk2Ctrl");
static_assert(kWsSk3CntOff + 2ull * kSk2MaxGroups * kSk3Warps * sizeof(int32_t) <= kWsCtrlBytes,
              "control area too small for the cooperative stream-K arrival counters");
size_t fixed_workspace_bytes(int batch, int num_q_heads, int head_size, int splits) {
  const size_t rows = static_cast<size_t>(batch) * num_q_heads * splits;
  return kWsCtrlBytes + rows * head_size * sizeof(float) + rows * sizeof(float) + 256;
}
// Workspace layout (both schedules): [0, 256) control words kept zero between launches (stream-K
// claim counter + finished-warp counter), then the partial O rows, the partial LSEs and (stream-K)
// the tiles-prefix plan. Stream-K has two partial slots per chunk: one static chunk per warp of the
// persistent grid plus at most kSkDynPerWarp dynamic chunks per warp.
// Share of the stream scheduled dynamically (B200ATTN_SK_DYN = shift, share = 1 / 2^shift). Off by
// default: measured on B200 (profiles/README.md) it evens out the warps' finish times but the kernel
// is HBM-bound throughout, so the gain (<= 2 %) is smaller than the extra partials cost the reduce.
int sk_dyn_shift() {
  const char* v = getenv("B200ATTN_SK_DYN");
  if (v == nullptr || *v == '\0') return kSkDynShiftDefault;
  const int x = atoi(v);
  return x < 0 ? 0 : (x > 8 ? 8 : x);
}
size_t sk_part_rows(int sms) {
  const int per_warp = 1 + (sk_dyn_shift() > 0 ? kSkDynPerWarp : 0);
  return static_cast<size_t>(2) * (static_cast<size_t>(sms) * kSkWarps * per_warp + 1) * kHeadsPerItem;
}
size_t sk_workspace_bytes(int batch, int sms) {
  const size_t rows = sk_part_rows(sms);
  return kWsCtrlBytes + rows * kHeadDim * sizeof(float) + rows * sizeof(float) +
         ((static_cast<size_t>(batch) + 1 + 3) / 4 * 4) * sizeof(int32_t) + 256;
}

// stream-K v2 (paged_decode_sk2.cu) workspace: [control | part_o | gpart_o | part_lse | gpart_lse | plan]
size_t sk2_part_rows(int nw) { return static_cast<size_t>(2) * (static_cast<size_t>(nw) + 1) * kHeadsPerItem; }
size_t sk2_workspace_bytes(int batch, int num_q_heads, int sms) {
  const int nw = sms * decode_sk2_warps();
  const size_t rows = sk2_part_rows(nw);
  // head-groups per sequence <= query heads (one per kv head and 8-head sub-group)
  return kWsCtrlBytes + rows * (kHeadDim + 1) * sizeof(float) +
         (sk2_plan_ints(batch, nw, num_q_heads) + 4) * sizeof(int32_t) + 256;
}
struct Sk2Layout {
  float *part_o, *part_lse;
  int32_t* plan;
  Sk2Ctrl* ctrl;
};
int sk2_calib_max() {  // weight updates before the SM weights are frozen (0 = uniform shares, never calibrate)
  const char* v = getenv("B200ATTN_SK2_CALIB");
  if (v == nullptr || *v == '\0') return kSk2CalibUpdatesDefault;
  const int x = atoi(v);
  return x < 0 ? 0 : x;
}
int sk3_calib_max() {
  const char* v = getenv("B200ATTN_SK3_CALIB");
  if (v == nullptr || *v == '\0') return kSk3CalibUpdatesDefault;
  const int x = atoi(v);
  return x < 0 ? 0 : x;
}
thread_local int g_sk2_parity = 0;
Sk2Layout sk2_layout(void* workspace, int sms) {
  const size_t rows = sk2_part_rows(sms * decode_sk2_warps());
  uint8_t* base = static_cast<uint8_t*>(workspace);
  Sk2Layout l;
  l.ctrl = reinterpret_cast<Sk2Ctrl*>(base + kWsSk2CtrlOff);
  l.part_o = reinterpret_cast<float*>(base + kWsCtrlBytes);
  l.part_lse = l.part_o + rows * kHeadDim;
  l.plan = reinterpret_cast<int32_t*>(l.part_lse + rows);
  return l;
}

// cooperative stream-K (paged_decode_sk3.cu) region, behind the stream-K v2 region:
// [partial rows: 2 parities x 2 * teams x 8 x 128 f32 | their LSEs | plan]
struct Sk3Layout {
  float *part_o, *part_lse;
  int32_t *plan, *cnt;
};
size_t sk3_region_off(int batch, int num_q_heads, int sms) {
  return (sk2_workspace_bytes(batch, num_q_heads, sms) + 255) & ~static_cast<size_t>(255);
}
size_t sk3_slot_rows(int sms) { return 2ull * 2ull * sms * decode_sk3_warps() * kHeadsPerItem; }
size_t sk3_workspace_bytes(int batch, int num_q_heads, int sms) {
  return sk3_region_off(batch, num_q_heads, sms) + sk3_slot_rows(sms) * (kHeadDim + 1) * sizeof(float) +
         (sk3_plan_ints(batch, sms) + 4) * sizeof(int32_t) + 256;
}
Sk3Layout sk3_layout(void* workspace, int batch, int num_q_heads, int sms) {
  uint8_t* base = static_cast<uint8_t*>(workspace);
  Sk3Layout l;
  l.cnt = reinterpret_cast<int32_t*>(base + kWsSk3CntOff);
  l.part_o = reinterpret_cast<float*>(base + sk3_region_off(batch, num_q_heads, sms));
  l.part_lse = l.part_o + sk3_slot_rows(sms) * kHeadDim;
  l.plan = reinterpret_cast<int32_t*>(l.part_lse + sk3_slot_rows(sms));
  return l;
}
thread_local int g_sk3_team = 0;  // debug: forced team size (0 = the plan kernel's rule)
int sk3_team_force() {
  if (g_sk3_team > 0) return g_sk3_team;
  const char* v = getenv("B200ATTN_SK3_TEAM");
  return (v == nullptr || *v == '\0') ? 0 : atoi(v);
}

thread_local unsigned long long* g_trace = nullptr;

constexpr int kMaxSplits = 128;
constexpr int kMinTilesPerSplit = 8;  // 128 tokens

int tiles_per_split_for(int max_seq_len, int num_splits) {
  int tiles = (max_seq_len + kTileTokens - 1) / kTileTokens;
  if (tiles < 1) tiles = 1;
  int c = (tiles + num_splits - 1) / num_splits;
  c = (c + 1) & ~1;  // whole 2-tile stages
  return c < 2 ? 2 : c;
}

}  // namespace

extern "C" {

int b200attn_abi_version(void) { return B200ATTN_ABI_VERSION; }
const char* b200attn_last_error(void) { return g_err; }
int b200attn_num_sms(void) { return cached_num_sms(); }

int b200attn_reshape_and_cache(const void* key, const void* value, int64_t key_stride_tok,
                               int64_t key_stride_head, int64_t value_stride_tok,
                               int64_t value_stride_head, void* k_cache, void* v_cache,
                               int64_t cache_stride_block, int64_t cache_stride_slot,
                               int64_t cache_stride_head, const int64_t* slot_mapping,
                               int32_t num_tokens, int32_t num_kv_heads, int32_t head_size,
                               int32_t block_size, int64_t num_blocks, int32_t kv_dtype,
                               const float* k_scale, const float* v_scale, int32_t scale_numel,
                               void* stream) {
  if (num_tokens < 0) return fail(B200ATTN_EINVAL, "num_tokens < 0");
  if (num_tokens == 0) return 0;
  if (key == nullptr || value == nullptr || slot_mapping == nullptr)
    return fail(B200ATTN_EINVAL, "null key/value/slot_mapping");
  if (int rc = check_cache_args(k_cache, v_cache, cache_stride_block, cache_stride_slot,
                                cache_stride_head, num_blocks, num_kv_heads, head_size, block_size,
                                kv_dtype))
    return rc;
  if ((key_stride_tok % 8) || (key_stride_head % 8) || (value_stride_tok % 8) ||
      (value_stride_head % 8) || (reinterpret_cast<uintptr_t>(key) % 16) ||
      (reinterpret_cast<uintptr_t>(value) % 16))
    return fail(B200ATTN_EINVAL, "key/value rows must be 16-byte aligned");
  const bool fp8 = kv_dtype == B200ATTN_KV_FP8_E4M3;
  if (fp8) {
    if (k_scale == nullptr || v_scale == nullptr) return fail(B200ATTN_EINVAL, "fp8 cache needs scales");
    if (scale_numel != 1 && scale_numel != num_kv_heads)
      return fail(B200ATTN_EINVAL, "scale_numel %d must be 1 or num_kv_heads", scale_numel);
  }
  const int sms = cached_num_sms();
  if (sms <= 0) return cuda_fail(cudaGetLastError(), "device query");
  CacheWriteParams p;
  p.key = static_cast<const __nv_bfloat16*>(key);
  p.value = static_cast<const __nv_bfloat16*>(value);
  p.key_stride_tok = key_stride_tok;
  p.key_stride_head = key_stride_head;
  p.value_stride_tok = value_stride_tok;
  p.value_stride_head = value_stride_head;
  p.k_cache = static_cast<uint8_t*>(k_cache);
  p.v_cache = static_cast<uint8_t*>(v_cache);
  p.cs_block = cache_stride_block;
  p.cs_slot = cache_stride_slot;
  p.cs_head = cache_stride_head;
  p.slot_mapping = slot_mapping;
  p.num_tokens = num_tokens;
  p.num_kv_heads = num_kv_heads;
  p.head_size = head_size;
  p.block_size = block_size;
  p.num_blocks = num_blocks;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.scale_stride = scale_numel > 1 ? 1 : 0;
  cudaError_t e = launch_cache_write(p, fp8, sms, env_flag("B200ATTN_PDL", true), static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "cache_write launch");
}

int b200attn_rope_and_cache(void* query, int64_t q_stride_tok, int32_t num_q_heads, void* key, const void* value,
                            int64_t key_stride_tok, int64_t key_stride_head, int64_t value_stride_tok,
                            int64_t value_stride_head, const int64_t* positions, int32_t pos_rows,
                            int64_t pos_stride_row, const void* cos_sin_cache, int64_t cos_sin_stride,
                            int64_t max_position, int32_t rot_dim, int32_t is_neox, const int32_t* mrope_section,
                            int32_t mrope_interleaved, void* k_cache, void* v_cache, int64_t cache_stride_block,
                            int64_t cache_stride_slot, int64_t cache_stride_head, const int64_t* slot_mapping,
                            int32_t num_tokens, int32_t num_slot_tokens, int32_t num_kv_heads, int32_t head_size,
                            int32_t block_size, int64_t num_blocks, int32_t kv_dtype, const float* k_scale,
                            const float* v_scale, int32_t scale_numel, void* stream) {
  (void)max_position;
  if (num_tokens < 0 || num_slot_tokens < 0 || num_slot_tokens > num_tokens)
    return fail(B200ATTN_EINVAL, "need 0 <= num_slot_tokens <= num_tokens");
  if (num_tokens == 0) return 0;
  if (query == nullptr || key == nullptr || value == nullptr || positions == nullptr || cos_sin_cache == nullptr ||
      (num_slot_tokens > 0 && slot_mapping == nullptr))
    return fail(B200ATTN_EINVAL, "null query/key/value/positions/cos_sin_cache/slot_mapping");
  if (int rc = check_cache_args(k_cache, v_cache, cache_stride_block, cache_stride_slot, cache_stride_head,
                                num_blocks, num_kv_heads, head_size, block_size, kv_dtype))
    return rc;
  if (num_q_heads <= 0) return fail(B200ATTN_EINVAL, "num_q_heads must be positive");
  if (rot_dim <= 0 || rot_dim > head_size || (rot_dim % 16) != 0)
    return fail(B200ATTN_EUNSUPPORTED, "rot_dim %d must be a multiple of 16 in (0, head_size]", rot_dim);
  if (pos_rows != 1 && pos_rows != 3) return fail(B200ATTN_EINVAL, "positions must be [T] or [3, T]");
  if (pos_rows == 3) {
    if (mrope_section == nullptr) return fail(B200ATTN_EINVAL, "M-RoPE positions need mrope_section");
    if (mrope_section[0] + mrope_section[1] + mrope_section[2] != rot_dim / 2)
      return fail(B200ATTN_EINVAL, "mrope_section must sum to rot_dim / 2");
  }
  if ((q_stride_tok % 8) || (key_stride_tok % 8) || (key_stride_head % 8) || (value_stride_tok % 8) ||
      (value_stride_head % 8) || (cos_sin_stride % 8) || (reinterpret_cast<uintptr_t>(query) % 16) ||
      (reinterpret_cast<uintptr_t>(key) % 16) || (reinterpret_cast<uintptr_t>(value) % 16) ||
      (reinterpret_cast<uintptr_t>(cos_sin_cache) % 16))
    return fail(B200ATTN_EINVAL, "query/key/value/cos_sin rows must be 16-byte aligned");
  const bool fp8 = kv_dtype == B200ATTN_KV_FP8_E4M3;
  if (fp8) {
    if (k_scale == nullptr || v_scale == nullptr) return fail(B200ATTN_EINVAL, "fp8 cache needs scales");
    if (scale_numel != 1 && scale_numel != num_kv_heads)
      return fail(B200ATTN_EINVAL, "scale_numel %d must be 1 or num_kv_heads", scale_numel);
  }
  const int sms = cached_num_sms();
  if (sms <= 0) return cuda_fail(cudaGetLastError(), "device query");
  RopeCacheParams p{};
  p.query = static_cast<__nv_bfloat16*>(query);
  p.q_stride_tok = q_stride_tok;
  p.key = static_cast<__nv_bfloat16*>(key);
  p.value = static_cast<const __nv_bfloat16*>(value);
  p.key_stride_tok = key_stride_tok;
  p.key_stride_head = key_stride_head;
  p.value_stride_tok = value_stride_tok;
  p.value_stride_head = value_stride_head;
  p.positions = positions;
  p.pos_stride_row = pos_stride_row;
  p.pos_rows = pos_rows;
  for (int i = 0; i < 3; ++i) p.mrope_section[i] = (pos_rows == 3) ? mrope_section[i] : 0;
  p.mrope_interleaved = mrope_interleaved;
  p.cos_sin = static_cast<const __nv_bfloat16*>(cos_sin_cache);
  p.cos_sin_stride = cos_sin_stride;
  p.rot_dim = rot_dim;
  p.is_neox = is_neox;
  p.k_cache = static_cast<uint8_t*>(k_cache);
  p.v_cache = static_cast<uint8_t*>(v_cache);
  p.cs_block = cache_stride_block;
  p.cs_slot = cache_stride_slot;
  p.cs_head = cache_stride_head;
  p.slot_mapping = slot_mapping;
  p.num_tokens = num_tokens;
  p.num_slot_tokens = num_slot_tokens;
  p.num_q_heads = num_q_heads;
  p.num_kv_heads = num_kv_heads;
  p.head_size = head_size;
  p.block_size = block_size;
  p.num_blocks = num_blocks;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.scale_stride = scale_numel > 1 ? 1 : 0;
  cudaError_t e = launch_rope_cache(p, fp8, sms, env_flag("B200ATTN_PDL", true), static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "rope_cache launch");
}

int32_t b200attn_decode_num_splits(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                                   int32_t max_seq_len, int32_t num_sms) {
  if (batch <= 0 || num_q_heads <= 0 || num_kv_heads <= 0 || num_sms <= 0) return 1;
  const int group = num_q_heads / num_kv_heads;
  const int subgroups = (group + kHeadsPerItem - 1) / kHeadsPerItem;
  const int64_t units = static_cast<int64_t>(batch) * num_kv_heads * subgroups;
  int tiles = (max_seq_len + kTileTokens - 1) / kTileTokens;
  if (tiles < 1) tiles = 1;
  int smax = tiles / kMinTilesPerSplit;
  if (smax < 1) smax = 1;
  if (smax > kMaxSplits) smax = kMaxSplits;
  // cost model of the persistent kernel: every CTA runs ceil(items / SMs) items of
  // tiles_per_split tiles, plus a fixed per-item epilogue worth ~2 tiles of streaming
  int best = 1;
  int64_t best_cost = -1;
  for (int s = 1; s <= smax; ++s) {
    const int c = tiles_per_split_for(max_seq_len, s);
    if (s > 1 && static_cast<int64_t>(s - 1) * c >= tiles) continue;  // last split would be empty
    const int64_t rounds = (units * s + num_sms - 1) / num_sms;
    const int64_t cost = rounds * (c + 2);
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      best = s;
    }
  }
  return best;
}

size_t b200attn_decode_workspace_bytes(int32_t max_batch, int32_t num_q_heads, int32_t head_size,
                                       int32_t max_splits) {
  if (max_batch <= 0 || num_q_heads <= 0 || head_size <= 0 || max_splits <= 0) return 0;
  const size_t fixed = fixed_workspace_bytes(max_batch, num_q_heads, head_size, max_splits);
  // the default (stream-K) schedule needs a batch-independent partial area; size for both so
  // that either schedule can run on a workspace of this size
  int sms = cached_num_sms();
  if (sms <= 0) sms = 148;
  const size_t sk = sk_workspace_bytes(max_batch, sms);
  const size_t sk2 = sk2_workspace_bytes(max_batch, num_q_heads, sms);
  const size_t sk3 = sk3_workspace_bytes(max_batch, num_q_heads, sms);
  size_t m = fixed > sk ? fixed : sk;
  m = m > sk2 ? m : sk2;
  return m > sk3 ? m : sk3;
}

static int decode_impl(bool run_partial, bool run_reduce, const void* q, int64_t q_stride_tok,
                       const void* k_cache, const void* v_cache, int64_t cache_stride_block,
                       int64_t cache_stride_slot, int64_t cache_stride_head, int64_t num_blocks,
                       void* out, int64_t out_stride_tok, float* lse, const int32_t* block_table,
                       int64_t block_table_stride, const int32_t* seq_lens, int32_t batch,
                       int32_t max_seq_len, int32_t num_q_heads, int32_t num_kv_heads,
                       int32_t head_size, int32_t block_size, float softmax_scale, int32_t kv_dtype,
                       const float* k_scale, const float* v_scale, int32_t scale_numel,
                       int32_t num_splits, int32_t variant, int32_t flags, void* workspace, size_t workspace_bytes,
                       void* stream_v) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  if (batch < 0) return fail(B200ATTN_EINVAL, "batch < 0");
  if (batch == 0) return 0;
  if (out == nullptr || seq_lens == nullptr) return fail(B200ATTN_EINVAL, "null out/seq_lens");
  if (num_q_heads <= 0 || num_kv_heads <= 0 || (num_q_heads % num_kv_heads) != 0)
    return fail(B200ATTN_EINVAL, "num_q_heads %d not a multiple of num_kv_heads %d", num_q_heads,
                num_kv_heads);
  if (head_size != kHeadDim)
    return fail(B200ATTN_EUNSUPPORTED, "head_size %d not built (128 only)", head_size);
  if ((out_stride_tok % 4) || (reinterpret_cast<uintptr_t>(out) % 8))
    return fail(B200ATTN_EINVAL, "out rows must be 8-byte aligned");
  if (max_seq_len < 1) max_seq_len = 1;
  const int sms = cached_num_sms();
  if (sms <= 0) return cuda_fail(cudaGetLastError(), "device query");

  if (num_splits < 0 || num_splits > kMaxSplits)
    return fail(B200ATTN_EINVAL, "num_splits %d out of [0, %d]", num_splits, kMaxSplits);
  if (variant < B200ATTN_VARIANT_AUTO || variant > B200ATTN_VARIANT_SK3)
    return fail(B200ATTN_EINVAL, "unknown variant %d", variant);
  if (flags & ~(B200ATTN_DECODE_PLANNED | B200ATTN_DECODE_EARLY_KV_READ | B200ATTN_DECODE_EARLY_Q_READ))
    return fail(B200ATTN_EINVAL, "unknown decode flags 0x%x", flags);
  // schedule: cooperative stream-K (teams of warps share an item, merges in shared memory, no reduce launch)
  // unless the caller asks for a fixed split count or another kernel
  bool use_sk3 = variant == B200ATTN_VARIANT_SK3 || (variant == B200ATTN_VARIANT_AUTO && num_splits == 0);
  // fp8 caches: the kernel is bound by instruction issue (e4m3 -> f16 conversions), not by HBM, so the schedule with
  // the leanest streaming loop wins — the first-generation stream-K (per-warp chunks + reduce launch): 183 us against
  // 203 us at cfg4 (profiles/README.md). AUTO takes it where it fits (batch <= 2048); B200ATTN_FP8_SCHEDULE=coop
  // keeps the cooperative kernel.
  if (use_sk3 && variant == B200ATTN_VARIANT_AUTO && kv_dtype == B200ATTN_KV_FP8_E4M3 && batch <= kSkMaxBatch &&
      workspace != nullptr && sk_workspace_bytes(batch, sms) <= workspace_bytes) {
    const char* v = getenv("B200ATTN_FP8_SCHEDULE");
    if (v == nullptr || *v == '\0' || strcmp(v, "coop") != 0) {
      const int tpp = block_size / kTileTokens;
      if (tpp > 0 && tpp * kTileTokens == block_size && (tpp & (tpp - 1)) == 0) {
        use_sk3 = false;
        variant = B200ATTN_VARIANT_STREAMK;
        flags &= ~B200ATTN_DECODE_PLANNED;  // this schedule rebuilds its plan in the kernel: the step's plan is not used
      }
    }
  }
  if (use_sk3) {
    const int tpp = block_size / kTileTokens;
    const bool fits = workspace != nullptr && tpp > 0 && tpp * kTileTokens == block_size && (tpp & (tpp - 1)) == 0 &&
                      sk3_workspace_bytes(batch, num_q_heads, sms) <= workspace_bytes && sms <= kSk2MaxGroups &&
                      (reinterpret_cast<uintptr_t>(workspace) % 16) == 0;
    if (!fits) {
      if (variant == B200ATTN_VARIANT_SK3 || (flags & B200ATTN_DECODE_PLANNED))
        return fail(B200ATTN_EWORKSPACE,
                    "stream-K needs block_size = 16 * 2^k and a 16-byte aligned workspace of %zu B (got %zu)",
                    sk3_workspace_bytes(batch, num_q_heads, sms), workspace_bytes);
      use_sk3 = false;  // AUTO: fall back to the fixed-split schedule
      if (variant == B200ATTN_VARIANT_AUTO) variant = B200ATTN_VARIANT_STREAM;
    }
  }
  if (use_sk3) variant = B200ATTN_VARIANT_SK3;
  // schedule: stream-K v2 (equal tile chunks per warp, no split count, merge inside the kernel) unless the
  // caller asks for a fixed split count or another kernel
  bool use_sk2 = variant == B200ATTN_VARIANT_SK2;
  if (use_sk2) {
    const int tpp = block_size / kTileTokens;
    const bool fits = workspace != nullptr && tpp > 0 && (tpp & (tpp - 1)) == 0 &&
                      sk2_workspace_bytes(batch, num_q_heads, sms) <= workspace_bytes && sms <= kSk2MaxGroups &&
                      (reinterpret_cast<uintptr_t>(workspace) % 16) == 0;
    if (!fits) {
      if (variant == B200ATTN_VARIANT_SK2 || (flags & B200ATTN_DECODE_PLANNED))
        return fail(B200ATTN_EWORKSPACE,
                    "stream-K needs block_size = 16 * 2^k and a 16-byte aligned workspace of %zu B (got %zu)",
                    sk2_workspace_bytes(batch, num_q_heads, sms), workspace_bytes);
      use_sk2 = false;  // AUTO: fall back to the fixed-split schedule
    }
  }
  if (use_sk2) variant = B200ATTN_VARIANT_SK2;
  if ((flags & B200ATTN_DECODE_PLANNED) && !use_sk2 && !use_sk3)
    return fail(B200ATTN_EINVAL, "B200ATTN_DECODE_PLANNED is only meaningful for the stream-K schedule");
  bool use_sk = variant == B200ATTN_VARIANT_STREAMK;
  if (use_sk) {
    const int tpp = block_size / kTileTokens;
    const bool fits = batch <= kSkMaxBatch && workspace != nullptr && tpp > 0 && (tpp & (tpp - 1)) == 0 &&
                      sk_workspace_bytes(batch, sms) <= workspace_bytes &&
                      (reinterpret_cast<uintptr_t>(workspace) % 16) == 0;
    if (!fits) {
      if (variant == B200ATTN_VARIANT_STREAMK)
        return fail(B200ATTN_EWORKSPACE,
                    "stream-K needs batch <= %d, block_size = 16 * 2^k and a 16-byte aligned workspace of %zu B (got %zu)",
                    kSkMaxBatch, sk_workspace_bytes(batch, sms), workspace_bytes);
      use_sk = false;  // AUTO: fall back to the fixed-split schedule
    }
  }
  if (variant == B200ATTN_VARIANT_AUTO || variant == B200ATTN_VARIANT_STREAMK)
    variant = use_sk ? B200ATTN_VARIANT_STREAMK : B200ATTN_VARIANT_STREAM;
  float* part_o = nullptr;
  float* part_lse = nullptr;
  int32_t* plan_cum = nullptr;
  Sk2Layout sk2{};
  Sk3Layout sk3{};
  if (use_sk3) {
    num_splits = 1;
    sk2 = sk2_layout(workspace, sms);  // the control words (claims, SM rates) are shared
    sk3 = sk3_layout(workspace, batch, num_q_heads, sms);
  } else if (use_sk2) {
    num_splits = 1;
    sk2 = sk2_layout(workspace, sms);
    part_o = sk2.part_o;
    part_lse = sk2.part_lse;
  } else if (use_sk) {
    num_splits = 1;
    const size_t rows = sk_part_rows(sms);
    part_o = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + kWsCtrlBytes);
    part_lse = part_o + rows * kHeadDim;
    plan_cum = reinterpret_cast<int32_t*>(part_lse + rows);
  } else {
    if (num_splits == 0) {
      num_splits = b200attn_decode_num_splits(batch, num_q_heads, num_kv_heads, max_seq_len, sms);
      // never ask for more splits than the caller's workspace can hold
      while (num_splits > 1 &&
             fixed_workspace_bytes(batch, num_q_heads, head_size, num_splits) > workspace_bytes)
        --num_splits;
    }
    if (num_splits > 1) {
      if (workspace == nullptr ||
          fixed_workspace_bytes(batch, num_q_heads, head_size, num_splits) > workspace_bytes)
        return fail(B200ATTN_EWORKSPACE, "workspace %zu B too small for %d splits (need %zu)",
                    workspace_bytes, num_splits,
                    fixed_workspace_bytes(batch, num_q_heads, head_size, num_splits));
      if (reinterpret_cast<uintptr_t>(workspace) % 16)
        return fail(B200ATTN_EINVAL, "workspace must be 16-byte aligned");
      const size_t rows = static_cast<size_t>(batch) * num_q_heads * num_splits;
      part_o = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + kWsCtrlBytes);
      part_lse = part_o + rows * head_size;
    }
  }
  const int tiles_per_split = tiles_per_split_for(max_seq_len, num_splits);

  DecodeParams p{};
  p.out = static_cast<__nv_bfloat16*>(out);
  p.out_stride_tok = out_stride_tok;
  p.lse = lse;
  p.seq_lens = seq_lens;
  p.batch = batch;
  p.num_q_heads = num_q_heads;
  p.num_kv_heads = num_kv_heads;
  p.group = num_q_heads / num_kv_heads;
  p.subgroups = (p.group + kHeadsPerItem - 1) / kHeadsPerItem;
  p.num_splits = num_splits;
  p.tiles_per_split = tiles_per_split;
  p.part_o = part_o;
  p.part_lse = part_lse;
  p.plan_cum = plan_cum;
  p.sk_ctrl = static_cast<int32_t*>(workspace);
  p.sk_dyn_shift = use_sk ? sk_dyn_shift() : 0;
  p.trace = g_trace;
  p.sk2_plan = sk2.plan;
  p.sk2_ctrl = sk2.ctrl;
  p.sk3_plan = sk3.plan;
  p.sk3_part_o = sk3.part_o;
  p.sk3_part_lse = sk3.part_lse;
  p.sk3_cnt = sk3.cnt;
  p.flags = 0;
  const int64_t items = static_cast<int64_t>(batch) * num_kv_heads * num_splits * p.subgroups;
  if (items > 0x7fffffff) return fail(B200ATTN_EINVAL, "too many work items");
  p.total_items = static_cast<int32_t>(items);

  if (run_partial) {
    if (q == nullptr || block_table == nullptr) return fail(B200ATTN_EINVAL, "null q/block_table");
    if (int rc = check_cache_args(k_cache, v_cache, cache_stride_block, cache_stride_slot,
                                  cache_stride_head, num_blocks, num_kv_heads, head_size, block_size,
                                  kv_dtype))
      return rc;
    if ((q_stride_tok % 2) || (reinterpret_cast<uintptr_t>(q) % 4))
      return fail(B200ATTN_EINVAL, "q rows must be 4-byte aligned");
    const bool fp8 = kv_dtype == B200ATTN_KV_FP8_E4M3;
    if (fp8) {
      if (variant != B200ATTN_VARIANT_STREAMK && variant != B200ATTN_VARIANT_SK2 && variant != B200ATTN_VARIANT_SK3)
        return fail(B200ATTN_EUNSUPPORTED,
                    "fp8 KV decode is built for the stream-K schedule only (variant AUTO/SK2/STREAMK, num_splits 0, "
                    "block_size = 16 * 2^k, workspace of b200attn_decode_workspace_bytes)");
      if (k_scale == nullptr || v_scale == nullptr) return fail(B200ATTN_EINVAL, "fp8 cache needs scales");
      if (scale_numel != 1 && scale_numel != num_kv_heads)
        return fail(B200ATTN_EINVAL, "scale_numel %d must be 1 or num_kv_heads", scale_numel);
    }
    p.q = static_cast<const __nv_bfloat16*>(q);
    p.q_stride_tok = q_stride_tok;
    p.k_cache = static_cast<const uint8_t*>(k_cache);
    p.v_cache = static_cast<const uint8_t*>(v_cache);
    p.cs_block = cache_stride_block;
    p.cs_slot = cache_stride_slot;
    p.cs_head = cache_stride_head;
    p.block_table = block_table;
    p.bt_stride = block_table_stride;
    p.block_size = block_size;
    p.tiles_per_page = block_size / kTileTokens;
    p.tpp_shift = -1;
    for (int sh = 0; sh < 16; ++sh)
      if ((1 << sh) == p.tiles_per_page) p.tpp_shift = sh;
    p.tpp_mask = p.tiles_per_page - 1;
    if ((variant == B200ATTN_VARIANT_STREAMK || variant == B200ATTN_VARIANT_SK2 || variant == B200ATTN_VARIANT_SK3) &&
        p.tpp_shift < 0) {
      // block sizes that are not 16 * 2^k go through the fixed-split kernel
      return fail(B200ATTN_EUNSUPPORTED, "stream-K needs block_size = 16 * 2^k (got %d)", block_size);
    }
    p.scale_log2 = softmax_scale * kLog2e;
    p.k_scale = k_scale;
    p.v_scale = v_scale;
    p.scale_stride = scale_numel > 1 ? 1 : 0;

    cudaError_t e;
    if (variant == B200ATTN_VARIANT_TMA) {
      const CUtensorMap *mk = nullptr, *mv = nullptr;
      CoordPerm pk, pv;
      if (int rc = get_kv_map(k_cache, cache_stride_block, cache_stride_slot, cache_stride_head,
                              num_blocks, num_kv_heads, block_size, head_size, fp8, &mk, &pk))
        return rc;
      // copy K's map before the second lookup can evict it from the direct-mapped cache
      const CUtensorMap tm_k = *mk;
      if (int rc = get_kv_map(v_cache, cache_stride_block, cache_stride_slot, cache_stride_head,
                              num_blocks, num_kv_heads, block_size, head_size, fp8, &mv, &pv))
        return rc;
      const int grid = static_cast<int>(items < sms ? items : sms);
      e = launch_decode_tma(tm_k, *mv, pk, p, grid, stream);
    } else if (variant == B200ATTN_VARIANT_CPASYNC) {
      e = launch_decode_cpasync(p, stream);
    } else if (variant == B200ATTN_VARIANT_STREAM) {
      e = launch_decode_stream(p, static_cast<int>(items < sms ? items : sms), stream);
    } else if (variant == B200ATTN_VARIANT_SK3) {
      const CUtensorMap *mk = nullptr, *mv = nullptr;
      if (int rc = get_kv_map5(k_cache, cache_stride_block, cache_stride_slot, cache_stride_head, num_blocks,
                               num_kv_heads, block_size, fp8, &mk))
        return rc;
      const CUtensorMap tm_k = *mk;  // copy before the second lookup can evict it
      if (int rc = get_kv_map5(v_cache, cache_stride_block, cache_stride_slot, cache_stride_head, num_blocks,
                               num_kv_heads, block_size, fp8, &mv))
        return rc;
      const bool pdl = env_flag("B200ATTN_PDL", true);
      if (!(flags & B200ATTN_DECODE_PLANNED)) {
        // no plan of this step in the workspace: make it now; the decode kernel then waits for it first
        if (cudaError_t pe = launch_sk3_plan(seq_lens, batch, num_kv_heads * p.subgroups, sms, block_table,
                                             block_table_stride, p.tpp_shift, sk3.plan, sk2.ctrl, sk3_calib_max(),
                                             sk3_team_force(), stream))
          return cuda_fail(pe, "decode plan launch");
        p.flags |= kSk2FlagWaitFirst;
      }
      if ((flags & B200ATTN_DECODE_EARLY_KV_READ) && env_flag("B200ATTN_EARLY_KV_READ", true))
        p.flags |= kSk2FlagEarlyKvRead;
      if ((flags & B200ATTN_DECODE_EARLY_Q_READ) && env_flag("B200ATTN_EARLY_Q_READ", true)) p.flags |= kSk3FlagEarlyQRead;
      // consecutive launches alternate their counter / slot sets and claim with growing ids; the step's plan launch
      // (which serialises the stream) resets the claim words, so the ids only have to grow inside a step
      static thread_local int launch_seq = 0;
      launch_seq = launch_seq >= 0x3fffffff ? 1 : launch_seq + 1;
      p.launch_id = launch_seq;
      e = launch_decode_sk3(tm_k, *mv, p, fp8, sms, pdl, stream);
    } else if (variant == B200ATTN_VARIANT_SK2) {
      const CUtensorMap *mk = nullptr, *mv = nullptr;
      if (int rc = get_kv_map5(k_cache, cache_stride_block, cache_stride_slot, cache_stride_head, num_blocks,
                               num_kv_heads, block_size, fp8, &mk))
        return rc;
      const CUtensorMap tm_k = *mk;  // copy before the second lookup can evict it
      if (int rc = get_kv_map5(v_cache, cache_stride_block, cache_stride_slot, cache_stride_head, num_blocks,
                               num_kv_heads, block_size, fp8, &mv))
        return rc;
      const bool pdl = env_flag("B200ATTN_PDL", true);
      if (!(flags & B200ATTN_DECODE_PLANNED)) {
        // no plan of this step in the workspace: make it now; the decode kernel then waits for it first
        if (cudaError_t pe = launch_sk2_plan(seq_lens, batch, num_kv_heads * p.subgroups, sms, block_table,
                                             block_table_stride, p.tiles_per_page, p.tpp_shift, sk2.plan, sk2.ctrl,
                                             sk2_calib_max(), stream))
          return cuda_fail(pe, "decode plan launch");
        p.flags |= kSk2FlagWaitFirst;
      }
      if ((flags & B200ATTN_DECODE_EARLY_KV_READ) && env_flag("B200ATTN_EARLY_KV_READ", true))
        p.flags |= kSk2FlagEarlyKvRead;
      if (env_flag("B200ATTN_SK2_DEBUG_INTERLEAVE", false)) p.flags |= kSk2FlagDebugInterleave;
      if ((g_sk2_parity ^= 1) != 0) p.flags |= kSk2FlagParity;  // consecutive launches alternate their claim words
      e = launch_decode_sk2(tm_k, *mv, p, fp8, sms, pdl, stream);
    } else if (variant == B200ATTN_VARIANT_STREAMK) {
      const CUtensorMap *mk = nullptr, *mv = nullptr;
      if (int rc = get_kv_map5(k_cache, cache_stride_block, cache_stride_slot, cache_stride_head, num_blocks,
                               num_kv_heads, block_size, fp8, &mk))
        return rc;
      const CUtensorMap tm_k = *mk;  // copy before the second lookup can evict it
      if (int rc = get_kv_map5(v_cache, cache_stride_block, cache_stride_slot, cache_stride_head, num_blocks,
                               num_kv_heads, block_size, fp8, &mv))
        return rc;
      e = launch_decode_streamk(tm_k, *mv, p, fp8, sms, env_flag("B200ATTN_PDL", true), stream);
    } else {
      return fail(B200ATTN_EINVAL, "unknown variant %d", variant);
    }
    if (e != cudaSuccess) return cuda_fail(e, "paged_decode partial launch");
  }
  if (run_reduce && use_sk2) {
    const bool pdl = run_partial && env_flag("B200ATTN_PDL", true);
    cudaError_t e = launch_sk2_reduce(p, sms, pdl, stream);
    if (e != cudaSuccess) return cuda_fail(e, "sk2_reduce launch");
  } else if (run_reduce && use_sk) {
    const bool pdl = run_partial && env_flag("B200ATTN_PDL", true);
    cudaError_t e = launch_streamk_reduce(p, sms, pdl, stream);
    if (e != cudaSuccess) return cuda_fail(e, "streamk_reduce launch");
  } else if (run_reduce && num_splits > 1) {
    const bool pdl = run_partial && env_flag("B200ATTN_PDL", true);
    cudaError_t e = launch_lse_reduce(static_cast<__nv_bfloat16*>(out), out_stride_tok, lse, seq_lens,
                                      batch, num_q_heads, num_splits, tiles_per_split, part_o,
                                      part_lse, pdl, stream);
    if (e != cudaSuccess) return cuda_fail(e, "lse_reduce launch");
  }
  return 0;
}

int b200attn_paged_decode(const void* q, int64_t q_stride_tok, const void* k_cache,
                          const void* v_cache, int64_t cache_stride_block, int64_t cache_stride_slot,
                          int64_t cache_stride_head, int64_t num_blocks, void* out,
                          int64_t out_stride_tok, float* lse, const int32_t* block_table,
                          int64_t block_table_stride, const int32_t* seq_lens, int32_t batch,
                          int32_t max_seq_len, int32_t num_q_heads, int32_t num_kv_heads,
                          int32_t head_size, int32_t block_size, float softmax_scale,
                          int32_t kv_dtype, const float* k_scale, const float* v_scale,
                          int32_t scale_numel, int32_t num_splits, int32_t variant, int32_t flags,
                          void* workspace, size_t workspace_bytes, void* stream) {
  return decode_impl(true, true, q, q_stride_tok, k_cache, v_cache, cache_stride_block,
                     cache_stride_slot, cache_stride_head, num_blocks, out, out_stride_tok, lse,
                     block_table, block_table_stride, seq_lens, batch, max_seq_len, num_q_heads,
                     num_kv_heads, head_size, block_size, softmax_scale, kv_dtype, k_scale, v_scale,
                     scale_numel, num_splits, variant, flags, workspace, workspace_bytes, stream);
}

int b200attn_paged_decode_partial(const void* q, int64_t q_stride_tok, const void* k_cache,
                                  const void* v_cache, int64_t cache_stride_block,
                                  int64_t cache_stride_slot, int64_t cache_stride_head,
                                  int64_t num_blocks, void* out, int64_t out_stride_tok, float* lse,
                                  const int32_t* block_table, int64_t block_table_stride,
                                  const int32_t* seq_lens, int32_t batch, int32_t max_seq_len,
                                  int32_t num_q_heads, int32_t num_kv_heads, int32_t head_size,
                                  int32_t block_size, float softmax_scale, int32_t kv_dtype,
                                  const float* k_scale, const float* v_scale, int32_t scale_numel,
                                  int32_t num_splits, int32_t variant, int32_t flags, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  if (variant != B200ATTN_VARIANT_STREAMK && variant != B200ATTN_VARIANT_SK2 && num_splits < 1)
    return fail(B200ATTN_EINVAL, "partial needs an explicit num_splits >= 1");
  if (variant == B200ATTN_VARIANT_AUTO) variant = B200ATTN_VARIANT_STREAM;
  return decode_impl(true, false, q, q_stride_tok, k_cache, v_cache, cache_stride_block,
                     cache_stride_slot, cache_stride_head, num_blocks, out, out_stride_tok, lse,
                     block_table, block_table_stride, seq_lens, batch, max_seq_len, num_q_heads,
                     num_kv_heads, head_size, block_size, softmax_scale, kv_dtype, k_scale, v_scale,
                     scale_numel, num_splits, variant, flags, workspace, workspace_bytes, stream);
}

int b200attn_lse_reduce(void* out, int64_t out_stride_tok, float* lse, const int32_t* seq_lens,
                        int32_t batch, int32_t max_seq_len, int32_t num_q_heads, int32_t num_kv_heads,
                        int32_t head_size, int32_t num_splits, int32_t variant, const void* workspace,
                        size_t workspace_bytes, void* stream) {
  if (variant != B200ATTN_VARIANT_STREAMK && num_splits < 1)
    return fail(B200ATTN_EINVAL, "reduce needs an explicit num_splits >= 1");
  if (variant == B200ATTN_VARIANT_AUTO) variant = B200ATTN_VARIANT_STREAM;
  return decode_impl(false, true, nullptr, 0, nullptr, nullptr, 0, 0, 0, 0, out, out_stride_tok, lse,
                     nullptr, 0, seq_lens, batch, max_seq_len, num_q_heads, num_kv_heads, head_size, 16,
                     1.f, B200ATTN_KV_BF16, nullptr, nullptr, 1, num_splits, variant, 0,
                     const_cast<void*>(workspace), workspace_bytes, stream);
}

// ---------------------------------------------------------------------------------------------
// Decode without a caller workspace and without any allocation: the split-KV partials live in a device array that
// is part of the library image. No counterpart in the reference; added for
// callers that cannot call cudaMalloc (graph capture, tight serving loops, harnesses that forbid it).
//
// In HOST code the array's name is only a stand-in for the device symbol — passing it to a kernel would hand the
// kernel a host address. Its device address is looked up once per device with cudaGetSymbolAddress (that device
// current) and cached below.
// ---------------------------------------------------------------------------------------------
constexpr size_t kStaticScratchBytes = B200ATTN_STATIC_SCRATCH_BYTES;
constexpr int kStaticMaxDevices = 64;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];
static_assert(kWsCtrlBytes + static_cast<size_t>(B200ATTN_STATIC_MAX_ROWS) * (kHeadDim + 1) * sizeof(float) + 256 <=
                  kStaticScratchBytes,
              "B200ATTN_STATIC_MAX_ROWS partial rows do not fit the static scratch array");
static void* g_static_scratch_ptr[kStaticMaxDevices];  // device address of g_static_scratch per device (0 = not asked yet)

static int static_scratch_address(void** ptr) {
  int dev = 0;
  if (cudaError_t e = cudaGetDevice(&dev)) return cuda_fail(e, "cudaGetDevice (static scratch)");
  if (dev < 0 || dev >= kStaticMaxDevices)
    return fail(B200ATTN_EUNSUPPORTED, "static scratch: device ordinal %d out of [0, %d)", dev, kStaticMaxDevices);
  void* p = __atomic_load_n(&g_static_scratch_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    // the same value whichever thread gets here first: no lock needed
    if (cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch))
      return cuda_fail(e, "cudaGetSymbolAddress (static scratch)");
    if (p == nullptr) return fail(B200ATTN_EDRIVER, "cudaGetSymbolAddress returned a null static scratch address");
    __atomic_store_n(&g_static_scratch_ptr[dev], p, __ATOMIC_RELEASE);
  }
  *ptr = p;
  return 0;
}

size_t b200attn_static_scratch_bytes(void) { return kStaticScratchBytes; }

int b200attn_paged_decode_static(const void* q, int64_t q_stride_tok, const void* k_cache, const void* v_cache,
                                 int64_t cache_stride_block, int64_t cache_stride_slot, int64_t cache_stride_head,
                                 int64_t num_blocks, void* out, int64_t out_stride_tok, float* lse,
                                 const int32_t* block_table, int64_t block_table_stride, const int32_t* seq_lens,
                                 int32_t batch, int32_t max_seq_len, int32_t num_q_heads, int32_t num_kv_heads,
                                 int32_t head_size, int32_t block_size, float softmax_scale, int32_t kv_dtype,
                                 int32_t num_splits, void* stream) {
  if (batch < 0) return fail(B200ATTN_EINVAL, "batch < 0");
  if (batch == 0) return 0;
  if (kv_dtype != B200ATTN_KV_BF16)
    return fail(B200ATTN_EUNSUPPORTED,
                "b200attn_paged_decode_static runs the fixed-split schedule, which is built for bf16 caches only");
  if (num_q_heads <= 0 || num_kv_heads <= 0 || (num_q_heads % num_kv_heads) != 0)
    return fail(B200ATTN_EINVAL, "num_q_heads %d not a multiple of num_kv_heads %d", num_q_heads, num_kv_heads);
  if (num_splits < 0 || num_splits > kMaxSplits)
    return fail(B200ATTN_EINVAL, "num_splits %d out of [0, %d]", num_splits, kMaxSplits);
  const int sms = cached_num_sms();
  if (sms <= 0) return cuda_fail(cudaGetLastError(), "device query");
  int splits = num_splits;
  if (splits == 0)
    splits = b200attn_decode_num_splits(batch, num_q_heads, num_kv_heads, max_seq_len < 1 ? 1 : max_seq_len, sms);
  // one KV range per sequence writes `out` directly and needs no scratch at all
  const int64_t rows = splits > 1 ? static_cast<int64_t>(batch) * num_q_heads * splits : 0;
  if (rows > B200ATTN_STATIC_MAX_ROWS)
    return fail(B200ATTN_EWORKSPACE,
                "static scratch holds %d partial rows (%zu B array), batch %d x %d query heads x %d splits needs %lld: "
                "pass a smaller num_splits (1 needs none) or use b200attn_paged_decode with a workspace of your own",
                B200ATTN_STATIC_MAX_ROWS, kStaticScratchBytes, batch, num_q_heads, splits,
                static_cast<long long>(rows));
  void* scratch = nullptr;
  if (int rc = static_scratch_address(&scratch)) return rc;
  // Nothing in the array has to start at zero for this schedule: the reduce launch reads exactly the partial rows
  // (sequence, head, split < splits of that sequence) that the decode launch of the same call wrote; what an
  // earlier call left in the other rows is never read.
  return decode_impl(true, true, q, q_stride_tok, k_cache, v_cache, cache_stride_block, cache_stride_slot,
                     cache_stride_head, num_blocks, out, out_stride_tok, lse, block_table, block_table_stride,
                     seq_lens, batch, max_seq_len, num_q_heads, num_kv_heads, head_size, block_size, softmax_scale,
                     kv_dtype, nullptr, nullptr, 1, splits, B200ATTN_VARIANT_STREAM, 0, scratch, kStaticScratchBytes,
                     stream);
}

int b200attn_decode_plan(const int32_t* seq_lens, int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                         const int32_t* block_table, int64_t block_table_stride, int32_t block_size,
                         void* workspace, size_t workspace_bytes, void* stream) {
  if (batch < 0) return fail(B200ATTN_EINVAL, "batch < 0");
  if (batch == 0) return 0;
  if (seq_lens == nullptr || block_table == nullptr) return fail(B200ATTN_EINVAL, "null seq_lens/block_table");
  if (num_q_heads <= 0 || num_kv_heads <= 0 || (num_q_heads % num_kv_heads) != 0)
    return fail(B200ATTN_EINVAL, "num_q_heads %d not a multiple of num_kv_heads %d", num_q_heads, num_kv_heads);
  const int sms = cached_num_sms();
  if (sms <= 0) return cuda_fail(cudaGetLastError(), "device query");
  const int tpp = block_size / kTileTokens;
  int shift = -1;
  for (int sh = 0; sh < 16; ++sh)
    if ((1 << sh) == tpp && tpp * kTileTokens == block_size) shift = sh;
  if (shift < 0) return fail(B200ATTN_EUNSUPPORTED, "stream-K needs block_size = 16 * 2^k (got %d)", block_size);
  if (workspace == nullptr || (reinterpret_cast<uintptr_t>(workspace) % 16) ||
      sk2_workspace_bytes(batch, num_q_heads, sms) > workspace_bytes || sms > kSk2MaxGroups)
    return fail(B200ATTN_EWORKSPACE, "decode plan needs a 16-byte aligned workspace of %zu B (got %zu)",
                sk2_workspace_bytes(batch, num_q_heads, sms), workspace_bytes);
  const int group = num_q_heads / num_kv_heads;
  const int subgroups = (group + kHeadsPerItem - 1) / kHeadsPerItem;
  const Sk2Layout l = sk2_layout(workspace, sms);
  if (sk3_workspace_bytes(batch, num_q_heads, sms) <= workspace_bytes) {
    const Sk3Layout l3 = sk3_layout(workspace, batch, num_q_heads, sms);
    if (cudaError_t e3 = launch_sk3_plan(seq_lens, batch, num_kv_heads * subgroups, sms, block_table, block_table_stride,
                                         shift, l3.plan, l.ctrl, sk3_calib_max(), sk3_team_force(),
                                         static_cast<cudaStream_t>(stream)))
      return cuda_fail(e3, "decode plan launch");
  }
  if (sk3_workspace_bytes(batch, num_q_heads, sms) <= workspace_bytes && !env_flag("B200ATTN_PLAN_SK2", false))
    return 0;  // the second-generation plan (40 us) is only made where that kernel is still compared
  cudaError_t e = launch_sk2_plan(seq_lens, batch, num_kv_heads * subgroups, sms, block_table, block_table_stride, tpp,
                                  shift, l.plan, l.ctrl, sk2_calib_max(), static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "decode plan launch");
}

int b200attn_workspace_init(void* workspace, size_t workspace_bytes, void* stream) {
  if (workspace == nullptr || workspace_bytes < kWsCtrlBytes)
    return fail(B200ATTN_EWORKSPACE, "workspace must hold at least %zu bytes", kWsCtrlBytes);
  cudaError_t e = cudaMemsetAsync(workspace, 0, kWsCtrlBytes, static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "workspace_init memset");
}

void b200attn_debug_sk_schedule(int64_t total_tiles, int32_t num_warps, int32_t dyn_shift, int64_t* out5) {
  const SkSchedule sc = sk_schedule(total_tiles, num_warps, dyn_shift);
  out5[0] = sc.S;
  out5[1] = sc.ts;
  out5[2] = sc.cd;
  out5[3] = sc.nd;
  out5[4] = sc.dyn ? 1 : 0;
}

void b200attn_debug_set_trace(void* device_buffer) { g_trace = static_cast<unsigned long long*>(device_buffer); }
void b200attn_debug_set_team(int32_t warps) { g_sk3_team = warps; }
uint64_t b200attn_debug_map_encodes(void) { return g_map_encodes; }
size_t b200attn_debug_plan_offset(int32_t batch, int32_t num_q_heads) {
  int sms = cached_num_sms();
  if (sms <= 0) sms = 148;
  uint8_t* base = nullptr;
  return static_cast<size_t>(reinterpret_cast<uint8_t*>(sk3_layout(base, batch, num_q_heads, sms).plan) - base);
}

int b200attn_paged_prefill(const void* q, int64_t q_stride_tok, const void* k_cache, const void* v_cache,
                           int64_t cache_stride_block, int64_t cache_stride_slot, int64_t cache_stride_head,
                           int64_t num_blocks, void* out, int64_t out_stride_tok, const int32_t* block_table,
                           int64_t block_table_stride, const int32_t* seq_lens, const int32_t* query_start_loc,
                           int32_t batch, int32_t num_tokens, int32_t max_query_len, int32_t num_q_heads,
                           int32_t num_kv_heads, int32_t head_size, int32_t block_size, float softmax_scale,
                           int32_t kv_dtype, const float* k_scale, const float* v_scale, int32_t scale_numel,
                           void* stream) {
  (void)max_query_len;
  if (batch < 0 || num_tokens < 0) return fail(B200ATTN_EINVAL, "batch/num_tokens < 0");
  if (batch == 0 || num_tokens == 0) return 0;
  if (q == nullptr || out == nullptr || block_table == nullptr || seq_lens == nullptr || query_start_loc == nullptr)
    return fail(B200ATTN_EINVAL, "null q/out/block_table/seq_lens/query_start_loc");
  if (num_q_heads <= 0 || num_kv_heads <= 0 || (num_q_heads % num_kv_heads) != 0)
    return fail(B200ATTN_EINVAL, "num_q_heads %d not a multiple of num_kv_heads %d", num_q_heads, num_kv_heads);
  if (int rc = check_cache_args(k_cache, v_cache, cache_stride_block, cache_stride_slot, cache_stride_head,
                                num_blocks, num_kv_heads, head_size, block_size, kv_dtype))
    return rc;
  if ((q_stride_tok % 2) || (reinterpret_cast<uintptr_t>(q) % 4))
    return fail(B200ATTN_EINVAL, "q rows must be 4-byte aligned");
  if ((out_stride_tok % 4) || (reinterpret_cast<uintptr_t>(out) % 8))
    return fail(B200ATTN_EINVAL, "out rows must be 8-byte aligned");
  const int tpp = block_size / kTileTokens;
  int shift = -1;
  for (int sh = 0; sh < 16; ++sh)
    if ((1 << sh) == tpp) shift = sh;
  if (shift < 0) return fail(B200ATTN_EUNSUPPORTED, "prefill needs block_size = 16 * 2^k (got %d)", block_size);
  const bool fp8 = kv_dtype == B200ATTN_KV_FP8_E4M3;
  if (fp8) {
    if (k_scale == nullptr || v_scale == nullptr) return fail(B200ATTN_EINVAL, "fp8 cache needs scales");
    if (scale_numel != 1 && scale_numel != num_kv_heads)
      return fail(B200ATTN_EINVAL, "scale_numel %d must be 1 or num_kv_heads", scale_numel);
  }
  PrefillParams p{};
  p.q = static_cast<const __nv_bfloat16*>(q);
  p.q_stride_tok = q_stride_tok;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.out_stride_tok = out_stride_tok;
  p.block_table = block_table;
  p.bt_stride = block_table_stride;
  p.seq_lens = seq_lens;
  p.query_start_loc = query_start_loc;
  p.batch = batch;
  p.num_tokens = num_tokens;
  p.num_q_heads = num_q_heads;
  p.num_kv_heads = num_kv_heads;
  p.group = num_q_heads / num_kv_heads;
  p.subgroups = (p.group + kHeadsPerItem - 1) / kHeadsPerItem;
  p.tpp_shift = shift;
  p.tpp_mask = tpp - 1;
  p.scale_log2 = softmax_scale * kLog2e;
  p.k_scale = fp8 ? k_scale : nullptr;  // bf16 caches carry no scales (the tcgen05 kernel applies any it is given)
  p.v_scale = fp8 ? v_scale : nullptr;
  p.scale_stride = scale_numel > 1 ? 1 : 0;
  // bf16 caches: the tcgen05 / TMEM kernel (B200ATTN_PREFILL_UMMA=0 selects the mma.sync kernel, which also serves fp8)
  const bool umma = !fp8 && env_flag("B200ATTN_PREFILL_UMMA", kPrefillUmmaDefault);
  const CUtensorMap *mk = nullptr, *mv = nullptr;
  if (int rc = get_kv_map5(k_cache, cache_stride_block, cache_stride_slot, cache_stride_head, num_blocks, num_kv_heads,
                           block_size, fp8, &mk, umma ? 1 : 2))
    return rc;
  const CUtensorMap tm_k = *mk;  // copy before the second lookup can evict it
  if (int rc = get_kv_map5(v_cache, cache_stride_block, cache_stride_slot, cache_stride_head, num_blocks, num_kv_heads,
                           block_size, fp8, &mv, umma ? 1 : 2))
    return rc;
  cudaError_t e = umma ? launch_paged_prefill_umma(tm_k, *mv, p, static_cast<cudaStream_t>(stream))
                       : launch_paged_prefill(tm_k, *mv, p, fp8, static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "paged_prefill launch");
}

// ---------------------------------------------------------------------------------------------
// fp8 prefill on the tcgen05 kernel: stage the step's e4m3 pages as bf16 (cache_ops.cu), then run
// paged_prefill_umma_kernel over the staged pages with the scales folded in.
// staging = [pages | cum: batch + 1 i32 | staged block table: batch x max_pages i32], each part 1 KiB aligned
// ---------------------------------------------------------------------------------------------
namespace {
struct StageLayout {
  size_t cum_off, bt_off, pages_off, page_bytes, total;
  int64_t capacity_pages;
  int max_pages;
};
StageLayout stage_layout(int batch, int max_seq_len, int64_t total_kv_tokens, int num_kv_heads, int head_size,
                         int block_size) {
  StageLayout l{};
  l.max_pages = (max_seq_len + block_size - 1) / block_size;
  if (l.max_pages < 1) l.max_pages = 1;
  int64_t cap = static_cast<int64_t>(batch) * l.max_pages;
  if (total_kv_tokens > 0) {  // sum over the sequences of ceil(S / bs) <= sum(S) / bs + batch
    const int64_t tight = total_kv_tokens / block_size + batch;
    if (tight < cap) cap = tight;
  }
  l.capacity_pages = cap;
  // pages first: their address (and with it the cached tensor map) does not move with the batch
  const size_t a = 1024;
  l.page_bytes = static_cast<size_t>(num_kv_heads) * 2 * block_size * head_size * sizeof(__nv_bfloat16);
  l.pages_off = 0;
  l.cum_off = (static_cast<size_t>(cap) * l.page_bytes + a - 1) / a * a;
  l.bt_off = l.cum_off + ((static_cast<size_t>(batch) + 1) * sizeof(int32_t) + a - 1) / a * a;
  l.total = l.bt_off + (static_cast<size_t>(batch) * l.max_pages * sizeof(int32_t) + a - 1) / a * a;
  return l;
}
}  // namespace

size_t b200attn_prefill_staging_bytes(int32_t batch, int32_t max_seq_len, int64_t total_kv_tokens,
                                      int32_t num_kv_heads, int32_t head_size, int32_t block_size) {
  if (batch <= 0 || max_seq_len <= 0 || num_kv_heads <= 0 || head_size <= 0 || block_size <= 0) return 0;
  return stage_layout(batch, max_seq_len, total_kv_tokens, num_kv_heads, head_size, block_size).total;
}

int b200attn_paged_prefill_fp8_staged(const void* q, int64_t q_stride_tok, const void* k_cache, const void* v_cache,
                                      int64_t cache_stride_block, int64_t cache_stride_slot, int64_t cache_stride_head,
                                      int64_t num_blocks, void* out, int64_t out_stride_tok, const int32_t* block_table,
                                      int64_t block_table_stride, const int32_t* seq_lens,
                                      const int32_t* query_start_loc, int32_t batch, int32_t num_tokens,
                                      int32_t max_query_len, int32_t max_seq_len, int64_t total_kv_tokens,
                                      int32_t num_q_heads, int32_t num_kv_heads, int32_t head_size, int32_t block_size,
                                      float softmax_scale, const float* k_scale, const float* v_scale,
                                      int32_t scale_numel, void* staging, size_t staging_bytes, void* stream) {
  (void)max_query_len;
  if (batch < 0 || num_tokens < 0) return fail(B200ATTN_EINVAL, "batch/num_tokens < 0");
  if (batch == 0 || num_tokens == 0) return 0;
  if (q == nullptr || out == nullptr || block_table == nullptr || seq_lens == nullptr || query_start_loc == nullptr)
    return fail(B200ATTN_EINVAL, "null q/out/block_table/seq_lens/query_start_loc");
  if (num_q_heads <= 0 || num_kv_heads <= 0 || (num_q_heads % num_kv_heads) != 0)
    return fail(B200ATTN_EINVAL, "num_q_heads %d not a multiple of num_kv_heads %d", num_q_heads, num_kv_heads);
  if (int rc = check_cache_args(k_cache, v_cache, cache_stride_block, cache_stride_slot, cache_stride_head,
                                num_blocks, num_kv_heads, head_size, block_size, B200ATTN_KV_FP8_E4M3))
    return rc;
  if ((q_stride_tok % 2) || (reinterpret_cast<uintptr_t>(q) % 4))
    return fail(B200ATTN_EINVAL, "q rows must be 4-byte aligned");
  if ((out_stride_tok % 4) || (reinterpret_cast<uintptr_t>(out) % 8))
    return fail(B200ATTN_EINVAL, "out rows must be 8-byte aligned");
  const int tpp = block_size / kTileTokens;
  int shift = -1;
  for (int sh = 0; sh < 16; ++sh)
    if ((1 << sh) == tpp) shift = sh;
  if (shift < 0) return fail(B200ATTN_EUNSUPPORTED, "prefill needs block_size = 16 * 2^k (got %d)", block_size);
  if (k_scale == nullptr || v_scale == nullptr) return fail(B200ATTN_EINVAL, "fp8 cache needs scales");
  if (scale_numel != 1 && scale_numel != num_kv_heads)
    return fail(B200ATTN_EINVAL, "scale_numel %d must be 1 or num_kv_heads", scale_numel);
  if (max_seq_len < 1) return fail(B200ATTN_EINVAL, "max_seq_len must bound the sequences of the step (got %d)", max_seq_len);
  const StageLayout l = stage_layout(batch, max_seq_len, total_kv_tokens, num_kv_heads, head_size, block_size);
  if (staging == nullptr || staging_bytes < l.total || (reinterpret_cast<uintptr_t>(staging) % 1024))
    return fail(B200ATTN_EWORKSPACE, "fp8 prefill staging needs %zu B, 1 KiB aligned (got %zu)", l.total, staging_bytes);
  if (l.capacity_pages > 0x7fffffff) return fail(B200ATTN_EINVAL, "too many staged pages");
  uint8_t* base = static_cast<uint8_t*>(staging);
  StageFp8Params sp{};
  sp.k_cache = static_cast<const uint8_t*>(k_cache);
  sp.v_cache = static_cast<const uint8_t*>(v_cache);
  sp.cs_block = cache_stride_block;
  sp.cs_slot = cache_stride_slot;
  sp.cs_head = cache_stride_head;
  sp.num_blocks = num_blocks;
  sp.block_table = block_table;
  sp.bt_stride = block_table_stride;
  sp.seq_lens = seq_lens;
  sp.batch = batch;
  sp.max_pages = l.max_pages;
  sp.num_kv_heads = num_kv_heads;
  sp.block_size = block_size;
  sp.capacity_pages = l.capacity_pages;
  sp.cum = reinterpret_cast<int32_t*>(base + l.cum_off);
  sp.stage_bt = reinterpret_cast<int32_t*>(base + l.bt_off);
  sp.pages = reinterpret_cast<__nv_bfloat16*>(base + l.pages_off);
  if (cudaError_t e = launch_stage_fp8(sp, static_cast<cudaStream_t>(stream))) return cuda_fail(e, "fp8 staging launch");

  PrefillParams p{};
  p.q = static_cast<const __nv_bfloat16*>(q);
  p.q_stride_tok = q_stride_tok;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.out_stride_tok = out_stride_tok;
  p.block_table = sp.stage_bt;
  p.bt_stride = l.max_pages;
  p.seq_lens = seq_lens;
  p.query_start_loc = query_start_loc;
  p.batch = batch;
  p.num_tokens = num_tokens;
  p.num_q_heads = num_q_heads;
  p.num_kv_heads = num_kv_heads;
  p.group = num_q_heads / num_kv_heads;
  p.subgroups = (p.group + kHeadsPerItem - 1) / kHeadsPerItem;
  p.tpp_shift = shift;
  p.tpp_mask = tpp - 1;
  p.scale_log2 = softmax_scale * kLog2e;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.scale_stride = scale_numel > 1 ? 1 : 0;
  // the staged pages as a bf16 cache in the product layout [page][Hkv][K|V][bs][d]
  const int64_t s_slot = head_size, s_head = 2ll * block_size * head_size, s_block = s_head * num_kv_heads;
  const __nv_bfloat16* sk = sp.pages;
  const __nv_bfloat16* sv = sp.pages + static_cast<int64_t>(block_size) * head_size;
  const CUtensorMap *mk = nullptr, *mv = nullptr;
  // page dimension of the map = what the caller's buffer could hold: one encode per staging buffer, not per step
  const int64_t map_pages = static_cast<int64_t>(staging_bytes / l.page_bytes);
  if (int rc = get_kv_map5(sk, s_block, s_slot, s_head, map_pages, num_kv_heads, block_size, false, &mk, 1))
    return rc;
  const CUtensorMap tm_k = *mk;  // copy before the second lookup can evict it
  if (int rc = get_kv_map5(sv, s_block, s_slot, s_head, map_pages, num_kv_heads, block_size, false, &mv, 1))
    return rc;
  cudaError_t e = launch_paged_prefill_umma(tm_k, *mv, p, static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "paged_prefill (staged fp8) launch");
}

int b200attn_gather_kv(const void* cache, int64_t cache_stride_block, int64_t cache_stride_slot,
                       int64_t cache_stride_head, int64_t num_blocks, const int32_t* block_ids,
                       int32_t num_block_ids, int32_t seq_len, int32_t num_kv_heads,
                       int32_t head_size, int32_t block_size, int32_t kv_dtype, void* dst,
                       void* stream) {
  if (seq_len < 0 || num_block_ids < 0) return fail(B200ATTN_EINVAL, "negative sizes");
  if (seq_len == 0) return 0;
  if (cache == nullptr || block_ids == nullptr || dst == nullptr)
    return fail(B200ATTN_EINVAL, "null pointer");
  const int eb = kv_dtype == B200ATTN_KV_FP8_E4M3 ? 1 : 2;
  if ((head_size * eb) % 16 || block_size <= 0 || num_kv_heads <= 0)
    return fail(B200ATTN_EINVAL, "rows must be multiples of 16 bytes");
  if (((cache_stride_slot * eb) % 16) || ((cache_stride_head * eb) % 16) ||
      ((cache_stride_block * eb) % 16) || (reinterpret_cast<uintptr_t>(cache) % 16) ||
      (reinterpret_cast<uintptr_t>(dst) % 16))
    return fail(B200ATTN_EINVAL, "cache/dst must be 16-byte aligned");
  if (static_cast<int64_t>(num_block_ids) * block_size < seq_len)
    return fail(B200ATTN_EINVAL, "seq_len %d exceeds %d blocks of %d", seq_len, num_block_ids,
                block_size);
  const int sms = cached_num_sms();
  if (sms <= 0) return cuda_fail(cudaGetLastError(), "device query");
  GatherParams p;
  p.cache = static_cast<const uint8_t*>(cache);
  p.cs_block = cache_stride_block;
  p.cs_slot = cache_stride_slot;
  p.cs_head = cache_stride_head;
  p.num_blocks = num_blocks;
  p.block_ids = block_ids;
  p.num_block_ids = num_block_ids;
  p.seq_len = seq_len;
  p.num_kv_heads = num_kv_heads;
  p.head_size = head_size;
  p.block_size = block_size;
  p.elem_bytes = eb;
  p.dst = static_cast<uint8_t*>(dst);
  cudaError_t e = launch_gather_kv(p, sms, static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "gather_kv launch");
}

// ---------------------------------------------------------------------------------------------
// peer-memory all-reduce
// ---------------------------------------------------------------------------------------------
static_assert(kPeerFlagBytes == B200ATTN_PEER_FLAG_BYTES && kPeerMaxRanks == B200ATTN_PEER_MAX_RANKS &&
                  (kPeerMaxBlocks * (kPeerMaxRanks + 1) + 8) * 4 <= kPeerFlagBytes, "flag region");

int32_t b200attn_peer_allreduce_algo(int64_t bytes, int32_t world, int32_t has_multicast, int64_t push_slot_bytes) {
  // PUSH (no flag round trips) while the (world-1) x bytes each rank posts stay latency-sized and fit a slot.
  // Otherwise bytes crossing NVLink per rank: ONESHOT (world-1) x bytes, TWOSHOT 2 x (world-1)/world x bytes + one
  // more flag barrier: the same at world 2, so ONESHOT there; the crossover sits near 256 KiB at 4 and 8 ranks.
  if (world <= 1) return B200ATTN_AR_ONESHOT;
  if (push_slot_bytes > 0 && env_flag("B200ATTN_AR_PUSH", true)) {
    if (bytes <= push_slot_bytes && (world - 1) * bytes <= (2 << 20)) return B200ATTN_AR_PUSH;
    // PUSH2 needs the longest slice (<= bytes/world + 16 * world) to fit half a slot
    if (world >= 4 && bytes / world + 16 * world <= push_slot_bytes / 2) return B200ATTN_AR_PUSH2;
  }
  if (world <= 2) return B200ATTN_AR_ONESHOT;
  if (has_multicast && env_flag("B200ATTN_AR_MULTIMEM", false) && bytes >= (64 << 10)) return B200ATTN_AR_MULTIMEM;
  return bytes <= (256 << 10) ? B200ATTN_AR_ONESHOT : B200ATTN_AR_TWOSHOT;
}

int b200attn_peer_allreduce(const void* const* peer_in, void* const* peer_scratch, void* const* peer_flags,
                            void* const* peer_push, int64_t push_slot_bytes, void* out, const void* mc_in,
                            void* mc_out, int64_t numel, int32_t world, int32_t rank, int32_t algo,
                            int32_t max_blocks, uint64_t timeout_ns, void* stream) {
  if (world != 1 && world != 2 && world != 4 && world != 8)
    return fail(B200ATTN_EUNSUPPORTED, "world %d (1, 2, 4 or 8 ranks)", world);
  if (rank < 0 || rank >= world) return fail(B200ATTN_EINVAL, "rank %d of %d", rank, world);
  if (numel < 0 || numel % 8) return fail(B200ATTN_EINVAL, "numel %lld must be a non-negative multiple of 8", (long long)numel);
  if (numel == 0) return 0;
  if (peer_in == nullptr || peer_flags == nullptr || out == nullptr) return fail(B200ATTN_EINVAL, "null pointer");
  if (push_slot_bytes < 0 || push_slot_bytes % 16) return fail(B200ATTN_EINVAL, "push_slot_bytes must be a multiple of 16");
  if (algo == B200ATTN_AR_AUTO)
    algo = b200attn_peer_allreduce_algo(numel * 2, world, mc_in != nullptr && mc_out != nullptr,
                                        peer_push != nullptr ? push_slot_bytes : 0);
  if (algo < B200ATTN_AR_ONESHOT || algo > B200ATTN_AR_PUSH2) return fail(B200ATTN_EINVAL, "algo %d", algo);
  if (algo == B200ATTN_AR_TWOSHOT && peer_scratch == nullptr) return fail(B200ATTN_EINVAL, "TWOSHOT needs peer_scratch");
  if (algo == B200ATTN_AR_MULTIMEM && (mc_in == nullptr || mc_out == nullptr))
    return fail(B200ATTN_EINVAL, "MULTIMEM needs multicast addresses");
  if (algo == B200ATTN_AR_PUSH && (peer_push == nullptr || numel * 2 > push_slot_bytes))
    return fail(B200ATTN_EINVAL, "PUSH needs peer_push with slots of at least %lld bytes", (long long)(numel * 2));
  const int64_t n_vec_all = numel / 8;
  const int64_t longest_slice = n_vec_all - (world - 1) * (n_vec_all / world);
  if (algo == B200ATTN_AR_PUSH2 && (peer_push == nullptr || world < 2 || longest_slice > push_slot_bytes / 32))
    return fail(B200ATTN_EINVAL, "PUSH2 needs >= 2 ranks and peer_push slots of at least %lld bytes",
                (long long)(longest_slice * 32));
  const bool pushy = algo == B200ATTN_AR_PUSH || algo == B200ATTN_AR_PUSH2;
  PeerArgs a;
  memset(&a, 0, sizeof(a));
  for (int r = 0; r < world; ++r) {
    const bool need_in = !pushy || r == rank;  // PUSH / PUSH2 read only the local input
    if (peer_flags[r] == nullptr || (reinterpret_cast<uintptr_t>(peer_flags[r]) % 16) ||
        (need_in && (peer_in[r] == nullptr || (reinterpret_cast<uintptr_t>(peer_in[r]) % 16))))
      return fail(B200ATTN_EINVAL, "peer_in/peer_flags[%d] null or not 16-byte aligned", r);
    a.in[r] = peer_in[r];
    a.ctl.flags[r] = static_cast<uint32_t*>(peer_flags[r]);
    if (algo == B200ATTN_AR_TWOSHOT) {
      if (peer_scratch[r] == nullptr || (reinterpret_cast<uintptr_t>(peer_scratch[r]) % 16))
        return fail(B200ATTN_EINVAL, "peer_scratch[%d] null or not 16-byte aligned", r);
      a.scratch[r] = peer_scratch[r];
    }
    if (pushy) {
      if (peer_push[r] == nullptr || (reinterpret_cast<uintptr_t>(peer_push[r]) % 16))
        return fail(B200ATTN_EINVAL, "peer_push[%d] null or not 16-byte aligned", r);
      a.push[r] = peer_push[r];
    }
  }
  if (reinterpret_cast<uintptr_t>(out) % 16) return fail(B200ATTN_EINVAL, "out must be 16-byte aligned");
  if (algo != B200ATTN_AR_MULTIMEM && out == peer_in[rank]) return fail(B200ATTN_EINVAL, "out may not alias the input region");
  a.out = out;
  a.mc_in = mc_in;
  a.mc_out = mc_out;
  a.push_slot_vec = push_slot_bytes / 16;
  a.ctl.world = world;
  a.ctl.rank = rank;
  a.ctl.timeout_ns = timeout_ns ? timeout_ns : 10000000000ull;
  a.ctl.soft_timeout = env_flag("B200ATTN_PEER_SOFT_TIMEOUT", false) ? 1 : 0;
  const int64_t n_vec = numel / 8;
  const bool whole = algo == B200ATTN_AR_ONESHOT || algo == B200ATTN_AR_PUSH;  // every rank touches every vector
  const int64_t per_rank = whole ? n_vec : n_vec - (world - 1) * (n_vec / world);
  if (max_blocks <= 0) max_blocks = 64;
  if (pushy && reinterpret_cast<uintptr_t>(out) == reinterpret_cast<uintptr_t>(peer_in[rank]))
    return fail(B200ATTN_EINVAL, "out may not alias the input");
  if (max_blocks > kPeerMaxBlocks) max_blocks = kPeerMaxBlocks;
  int64_t grid = (per_rank + kPeerThreads - 1) / kPeerThreads;
  if (grid < 1) grid = 1;
  if (grid > max_blocks) grid = max_blocks;
  cudaError_t e = launch_peer_allreduce(a, n_vec, algo, static_cast<int>(grid), env_flag("B200ATTN_PDL", true),
                                        static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "peer_allreduce launch");
}

int b200attn_peer_exchange(void* const* peer_rows, void* const* peer_flags, void* local_rows, int32_t group_size,
                           int32_t group_rank, int32_t first_row, int32_t num_rows, int32_t rows_per_rank,
                           int64_t row_bytes, int32_t scatter, uint64_t timeout_ns, void* stream) {
  if (group_size < 1 || group_size > kPeerMaxRanks) return fail(B200ATTN_EUNSUPPORTED, "group of %d ranks (1..8)", group_size);
  if (group_rank < 0 || group_rank >= group_size) return fail(B200ATTN_EINVAL, "rank %d of %d", group_rank, group_size);
  if (peer_rows == nullptr || peer_flags == nullptr || local_rows == nullptr) return fail(B200ATTN_EINVAL, "null pointer");
  if (row_bytes <= 0 || row_bytes % 16 || row_bytes / 16 > 0x7fffffff) return fail(B200ATTN_EINVAL, "row_bytes must be a positive multiple of 16");
  if (first_row < 0 || num_rows < 0 || rows_per_rank < num_rows) return fail(B200ATTN_EINVAL, "bad row range");
  if (reinterpret_cast<uintptr_t>(local_rows) % 16) return fail(B200ATTN_EINVAL, "local_rows must be 16-byte aligned");
  PeerExchangeArgs a;
  memset(&a, 0, sizeof(a));
  for (int r = 0; r < group_size; ++r) {
    if (peer_rows[r] == nullptr || peer_flags[r] == nullptr || (reinterpret_cast<uintptr_t>(peer_rows[r]) % 16) ||
        (reinterpret_cast<uintptr_t>(peer_flags[r]) % 16))
      return fail(B200ATTN_EINVAL, "peer_rows/peer_flags[%d] null or not 16-byte aligned", r);
    a.peer_rows[r] = peer_rows[r];
    a.ctl.flags[r] = static_cast<uint32_t*>(peer_flags[r]);
  }
  a.ctl.world = group_size;
  a.ctl.rank = group_rank;
  a.ctl.timeout_ns = timeout_ns ? timeout_ns : 10000000000ull;
  a.ctl.soft_timeout = env_flag("B200ATTN_PEER_SOFT_TIMEOUT", false) ? 1 : 0;
  a.local_rows = local_rows;
  a.b0 = first_row;
  a.nrows = num_rows;
  a.row_vec = static_cast<int32_t>(row_bytes / 16);
  // every rank of the group must launch the same grid (the barrier pairs same-index blocks): size it from
  // rows_per_rank, which is the same everywhere, not from this rank's possibly shorter range
  int64_t grid = (static_cast<int64_t>(rows_per_rank) * group_size * a.row_vec + kPeerThreads - 1) / kPeerThreads;
  if (grid < 1) grid = 1;
  if (grid > 64) grid = 64;
  cudaError_t e = launch_peer_exchange(a, scatter != 0, static_cast<int>(grid), env_flag("B200ATTN_PDL", true),
                                       static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "peer_exchange launch");
}

int b200attn_peer_alloc(size_t bytes, void** ptr, void* handle64) {
  if (ptr == nullptr || handle64 == nullptr || bytes == 0) return fail(B200ATTN_EINVAL, "null pointer / zero size");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc (peer_alloc)");
  if ((e = cudaMemset(p, 0, bytes)) != cudaSuccess || (e = cudaDeviceSynchronize()) != cudaSuccess) {
    cudaFree(p);
    return cuda_fail(e, "cudaMemset (peer_alloc)");
  }
  cudaIpcMemHandle_t h;
  if ((e = cudaIpcGetMemHandle(&h, p)) != cudaSuccess) {
    cudaFree(p);
    return cuda_fail(e, "cudaIpcGetMemHandle");
  }
  memcpy(handle64, &h, sizeof(h));
  *ptr = p;
  return 0;
}

int b200attn_peer_open(const void* handle64, void** ptr) {
  if (ptr == nullptr || handle64 == nullptr) return fail(B200ATTN_EINVAL, "null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return cuda_fail(e, "cudaIpcOpenMemHandle");
  *ptr = p;
  return 0;
}

int b200attn_peer_close(void* ptr) {
  if (ptr == nullptr) return 0;
  cudaError_t e = cudaIpcCloseMemHandle(ptr);
  return e == cudaSuccess ? 0 : cuda_fail(e, "cudaIpcCloseMemHandle");
}

int b200attn_peer_free(void* ptr) {
  if (ptr == nullptr) return 0;
  cudaError_t e = cudaFree(ptr);
  return e == cudaSuccess ? 0 : cuda_fail(e, "cudaFree (peer_free)");
}

}  // extern "C"
