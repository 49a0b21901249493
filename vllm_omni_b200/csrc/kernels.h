// Internal launcher declarations shared by api.cu and the kernel translation units.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "decode_common.cuh"

namespace b200 {

// coordinate order of the 4-D KV tensor maps: which of (slot, head, block) is tensor-map
// dimension 1, 2, 3 (dimension 0 is always head_size). Values: 0 = slot, 1 = head, 2 = block.
struct CoordPerm {
  int32_t dim1, dim2, dim3;
};

cudaError_t launch_decode_tma(const CUtensorMap& tm_k, const CUtensorMap& tm_v, CoordPerm perm,
                              const DecodeParams& p, int grid, cudaStream_t stream);
cudaError_t launch_decode_cpasync(const DecodeParams& p, cudaStream_t stream);
cudaError_t launch_decode_stream(const DecodeParams& p, int grid, cudaStream_t stream);
size_t decode_stream_smem_bytes();
cudaError_t launch_decode_streamk(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const DecodeParams& p, bool fp8,
                                  int grid, bool pdl, cudaStream_t stream);
cudaError_t launch_streamk_reduce(const DecodeParams& p, int grid_decode, bool pdl, cudaStream_t stream);
size_t decode_streamk_smem_bytes(bool fp8);
// stream-K v2: plan once per step, rate-aware chunk groups, plan-driven reduce kernel (paged_decode_sk2.cu)
cudaError_t launch_sk2_plan(const int32_t* seq_lens, int batch, int hs_count, int groups, const int32_t* block_table,
                            int64_t bt_stride, int tiles_per_page, int tpp_shift, int32_t* plan, Sk2Ctrl* ctrl,
                            int calib_max, cudaStream_t stream);
size_t sk2_ctrl_bytes();
cudaError_t launch_decode_sk2(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const DecodeParams& p, bool fp8,
                              int grid, bool pdl, cudaStream_t stream);
size_t decode_sk2_smem_bytes(bool fp8);
int decode_sk2_warps();
size_t sk2_plan_ints(int batch, int nw, int hs_count);
cudaError_t launch_sk2_reduce(const DecodeParams& p, int grid_decode, bool pdl, cudaStream_t stream);
// cooperative stream-K (paged_decode_sk3.cu): teams of warps share an item, merges in shared memory, no reduce launch
cudaError_t launch_sk3_plan(const int32_t* seq_lens, int batch, int hs_count, int groups, const int32_t* block_table,
                            int64_t bt_stride, int tpp_shift, int32_t* plan, Sk2Ctrl* ctrl, int calib_max, int team_force,
                            cudaStream_t stream);
cudaError_t launch_decode_sk3(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const DecodeParams& p, bool fp8,
                              int grid, bool pdl, cudaStream_t stream);
size_t decode_sk3_smem_bytes(bool fp8);
size_t sk3_plan_ints(int batch, int groups);
int decode_sk3_warps();
size_t decode_tma_smem_bytes();
size_t decode_cpasync_smem_bytes();

cudaError_t launch_lse_reduce(__nv_bfloat16* out, int64_t out_stride_tok, float* lse,
                              const int32_t* seq_lens, int batch, int num_q_heads, int num_splits,
                              int tiles_per_split, const float* part_o, const float* part_lse,
                              bool pdl, cudaStream_t stream);

struct PrefillParams {
  const __nv_bfloat16* q;
  int64_t q_stride_tok;
  __nv_bfloat16* out;
  int64_t out_stride_tok;
  const int32_t* block_table;
  int64_t bt_stride;
  const int32_t* seq_lens;
  const int32_t* query_start_loc;
  int32_t batch, num_tokens, num_q_heads, num_kv_heads;
  int32_t group, subgroups;
  int32_t tpp_shift, tpp_mask;  // 16-token tiles per page = 2^tpp_shift
  float scale_log2;
  const float* k_scale;
  const float* v_scale;
  int32_t scale_stride;
};
cudaError_t launch_paged_prefill(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const PrefillParams& p, bool fp8,
                                 cudaStream_t stream);
size_t prefill_smem_bytes(bool fp8);
cudaError_t launch_paged_prefill_umma(const CUtensorMap& tm_k, const CUtensorMap& tm_v, const PrefillParams& p,
                                      cudaStream_t stream);
size_t prefill_umma_smem_bytes();

// fp8 prefill staging (cache_ops.cu): e4m3 pages of the step's prefill sequences -> bf16 pages + their block table
struct StageFp8Params {
  const uint8_t* k_cache;
  const uint8_t* v_cache;
  int64_t cs_block, cs_slot, cs_head;  // bytes (= e4m3 elements)
  int64_t num_blocks;
  const int32_t* block_table;
  int64_t bt_stride;
  const int32_t* seq_lens;
  int32_t batch, max_pages, num_kv_heads, block_size;
  int64_t capacity_pages;
  int32_t* cum;          // [batch + 1] exclusive prefix of pages per sequence
  int32_t* stage_bt;     // [batch][max_pages] staged page ids
  __nv_bfloat16* pages;  // [capacity_pages][Hkv][2][bs][128]
};
cudaError_t launch_stage_fp8(const StageFp8Params& p, cudaStream_t stream);

struct CacheWriteParams {
  const __nv_bfloat16* key;
  const __nv_bfloat16* value;
  int64_t key_stride_tok, key_stride_head, value_stride_tok, value_stride_head;
  uint8_t* k_cache;
  uint8_t* v_cache;
  int64_t cs_block, cs_slot, cs_head;
  const int64_t* slot_mapping;
  int32_t num_tokens, num_kv_heads, head_size, block_size;
  int64_t num_blocks;
  const float* k_scale;
  const float* v_scale;
  int32_t scale_stride;
};
cudaError_t launch_cache_write(const CacheWriteParams& p, bool fp8, int num_sms, bool pdl, cudaStream_t stream);

struct RopeCacheParams {
  __nv_bfloat16* query;  // [T, Hq, d], rotated in place
  int64_t q_stride_tok;
  __nv_bfloat16* key;    // [T, Hkv, d], rotated in place
  const __nv_bfloat16* value;
  int64_t key_stride_tok, key_stride_head, value_stride_tok, value_stride_head;
  const int64_t* positions;  // [T] or [3, T]
  int64_t pos_stride_row;
  int32_t pos_rows;          // 1 = plain RoPE, 3 = M-RoPE
  int32_t mrope_section[3];
  int32_t mrope_interleaved;
  const __nv_bfloat16* cos_sin;  // [max_pos, rot_dim]: cos | sin
  int64_t cos_sin_stride;
  int32_t rot_dim, is_neox;
  uint8_t* k_cache;
  uint8_t* v_cache;
  int64_t cs_block, cs_slot, cs_head;
  const int64_t* slot_mapping;
  int32_t num_tokens, num_slot_tokens, num_q_heads, num_kv_heads, head_size, block_size;
  int64_t num_blocks;
  const float* k_scale;
  const float* v_scale;
  int32_t scale_stride;
};
cudaError_t launch_rope_cache(const RopeCacheParams& p, bool fp8, int num_sms, bool pdl, cudaStream_t stream);

struct GatherParams {
  const uint8_t* cache;
  int64_t cs_block, cs_slot, cs_head;  // elements
  int64_t num_blocks;
  const int32_t* block_ids;
  int32_t num_block_ids, seq_len, num_kv_heads, head_size, block_size, elem_bytes;
  uint8_t* dst;
};
cudaError_t launch_gather_kv(const GatherParams& p, int num_sms, cudaStream_t stream);

// ---- peer-memory all-reduce (peer_allreduce.cu) ----
constexpr int kPeerMaxRanks = 8;
constexpr int kPeerMaxBlocks = 128;   // flag slots exist for this many blocks per launch
constexpr int kPeerThreads = 512;
constexpr size_t kPeerFlagBytes = 8192;  // >= (kPeerMaxBlocks * (kPeerMaxRanks + 1) + 8) * 4, 16-byte multiple
constexpr int kPeerOneShot = 1, kPeerTwoShot = 2, kPeerMultimem = 3, kPeerPush = 4, kPeerPush2 = 5;
struct PeerCtl {                   // who takes part in a flag barrier
  uint32_t* flags[kPeerMaxRanks];  // every rank's flag region (kPeerFlagBytes, zeroed once), index = rank
  int32_t world, rank;
  uint64_t timeout_ns;
  int32_t soft_timeout;            // B200ATTN_PEER_SOFT_TIMEOUT=1 (tests): record + give up instead of trapping
};
struct PeerArgs {
  PeerCtl ctl;
  const void* in[kPeerMaxRanks];   // every rank's input region (peer-mapped), index = rank
  void* scratch[kPeerMaxRanks];    // every rank's scratch region (two-shot)
  void* out;                       // local result
  const void* mc_in;               // multicast view of the input region (multimem) or null
  void* mc_out;                    // multicast view of the output region (multimem) or null
  void* push[kPeerMaxRanks];       // every rank's receive region of the PUSH schedule: [3][world][push_slot_vec] x 16 B
  int64_t push_slot_vec;
};
// Row exchange inside a KV-head replica group (tp.ReplicaSplit): rank g of the group owns rows
// [B, row_bytes] (its query heads of every sequence); rank i serves sequences [b0_i, b0_i + per).
//   gather : local[(s - b0) * R + g] = peer_g[s]   for s in my sequence range   (pull, barrier first)
//   scatter: peer_g[s] = local[(s - b0) * R + g]                                  (push, barrier last)
struct PeerExchangeArgs {
  PeerCtl ctl;
  void* peer_rows[kPeerMaxRanks];  // every group rank's [B, row_bytes] region (peer-mapped)
  void* local_rows;                // [per, R, row_bytes] local staging (q_grp / out_grp)
  int32_t b0, nrows;               // my sequence range [b0, b0 + nrows)
  int32_t row_vec;                 // row_bytes / 16
};
cudaError_t launch_peer_exchange(const PeerExchangeArgs& a, bool scatter, int grid, bool pdl, cudaStream_t stream);
cudaError_t launch_peer_allreduce(const PeerArgs& a, int64_t n_vec, int algo, int grid, bool pdl, cudaStream_t stream);

}  // namespace b200
