/*
 * b200attn.h — C ABI of the B200 (sm_100a) paged-attention decode path.
 *
 * This is the drop-in boundary underneath the Python shim that implements vLLM's
 * AttentionBackend / AttentionImpl interface for vllm-omni's thinker/talker stages.
 * There is no pre-existing FFI on this path in the reference (it is a Python class
 * interface); every entry point below cites the reference/vLLM call it replaces.
 *
 * Conventions
 *   - plain C symbols; all pointers are DEVICE pointers unless marked "host";
 *   - strides are in ELEMENTS of the tensor's dtype (cache strides: elements of the
 *     cache dtype, i.e. bf16 or u8);
 *   - the caller owns every buffer; nothing is allocated, freed or synchronised here;
 *   - every function enqueues on `stream` (a cudaStream_t passed as void*) and returns;
 *   - return 0 = OK, negative = argument rejected on the host before any launch
 *     (see B200ATTN_E*), positive = cudaError_t of the failed launch;
 *   - no exceptions, no exit(); b200attn_last_error() gives a thread-local message;
 *   - safe under CUDA-graph capture: no host sync, no allocation, grids depend only
 *     on host arguments (sequence lengths are read on the device).
 *
 * Logical KV-cache layout (vLLM, `FlashAttentionBackend.get_kv_cache_shape`,
 * vllm/v1/attention/backends/flash_attn.py; reader constraint in the reference:
 * vllm_omni/distributed/omni_connectors/utils/kv_utils.py:417-450):
 *     kv_cache[2, num_blocks, block_size, num_kv_heads, head_size]
 * k_cache = kv_cache[0], v_cache = kv_cache[1]; the physical order is free and is
 * described by (stride_block, stride_slot, stride_head); head_size is contiguous.
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ATTN_ABI_VERSION 10

#if defined(__GNUC__)
#define B200ATTN_API __attribute__((visibility("default")))
#else
#define B200ATTN_API
#endif

/* kv_dtype */
#define B200ATTN_KV_BF16 0
#define B200ATTN_KV_FP8_E4M3 1

/* host-side argument errors (returned negated) */
#define B200ATTN_EINVAL 1      /* bad pointer / size / stride / alignment          */
#define B200ATTN_EUNSUPPORTED 2 /* shape or dtype this build has no kernel for       */
#define B200ATTN_EWORKSPACE 3  /* workspace too small for the requested split count */
#define B200ATTN_EDRIVER 4     /* CUDA driver entry point (tensor-map encode) failed */

/* decode kernel variants (for the ncu sweeps; 0 is what the backend uses) */
#define B200ATTN_VARIANT_AUTO 0
#define B200ATTN_VARIANT_TMA 1      /* persistent, TMA-staged pages, mbarrier ring       */
#define B200ATTN_VARIANT_CPASYNC 2  /* one CTA per work item, cp.async-staged pages      */
#define B200ATTN_VARIANT_STREAM 3   /* persistent, per-warp cp.async rings across items  */
#define B200ATTN_VARIANT_STREAMK 4  /* persistent, equal tile chunks per warp + reduce launch (round-1 default) */
#define B200ATTN_VARIANT_SK2 5      /* the same schedule planned once per step, chunk sizes follow the SMs' measured
                                       rates, plan-driven reduce launch (round 2 comparison arm) */
#define B200ATTN_VARIANT_SK3 6      /* cooperative stream-K (default): teams of warps walk an item together and fold
                                       their states in shared memory; items cut by a team boundary are folded by
                                       their owner through an arrival counter — no reduce launch */

/* b200attn_paged_decode flags */
#define B200ATTN_DECODE_PLANNED 1        /* b200attn_decode_plan was run on this workspace for exactly these
                                            seq_lens / block_table / batch / head counts / block_size (once per
                                            step; every layer of the step then skips the plan launch). "Once per
                                            step" also under CUDA-graph replay: a replayed step needs its plan
                                            launch (in the graph or before it) — the plan re-arms the words the
                                            step's launches claim their share of the stream with; launches
                                            replayed without it still compute the right result but lose the
                                            SM-aware shares. */
#define B200ATTN_DECODE_EARLY_KV_READ 2  /* the caller guarantees that, among this library's kernels still in
                                            flight when the decode kernel starts (they release their dependents
                                            early: programmatic dependent launch), the only KV-cache bytes being
                                            written are the new token's slot of each decoded sequence, i.e. its
                                            LAST 16-token tile — what a decode step's cache write does
                                            (gpu_ar_model_runner.py:421-456 slot mappings). The kernel may then
                                            stage all other KV tiles before griddepcontrol.wait. */
#define B200ATTN_DECODE_EARLY_Q_READ 4   /* the caller guarantees that `q` was complete before the kernel this launch
                                            follows was started (e.g. q comes out of a RoPE kernel, then this
                                            library's cache write, then this call): the decode kernel may then load
                                            q before griddepcontrol.wait. NOT valid after the fused RoPE + cache
                                            write (b200attn_rope_and_cache rotates q in place). */

B200ATTN_API int b200attn_abi_version(void);
B200ATTN_API const char* b200attn_last_error(void); /* host string, thread-local */

/* Number of SMs of the current device (cached); <=0 on error. */
B200ATTN_API int b200attn_num_sms(void);

/*
 * KV-cache write. Replaces vLLM `reshape_and_cache_flash`
 * (vllm/_custom_ops.py:2591-2610, called from
 * FlashAttentionImpl.do_kv_cache_update, vllm/v1/attention/backends/flash_attn.py:850-883;
 * semantics restated in vllm/v1/attention/ops/triton_reshape_and_cache_flash.py:44-124).
 *
 * For every token t < num_tokens with slot = slot_mapping[t] >= 0:
 *     block = slot / block_size, off = slot % block_size
 *     k_cache[block, off, h, :] = key[t, h, :]      (bf16: bit copy)
 *     k_cache[block, off, h, :] = e4m3_satfinite(float(key[t,h,:]) / k_scale[h or 0])
 * and the same for value / v_cache / v_scale. slot < 0 => token skipped.
 * key/value: bf16 [>=num_tokens, num_kv_heads, head_size], token and head strides given
 * (they are slices of the fused qkv projection output in the reference's model code,
 * vllm_omni/model_executor/models/qwen2_5_omni/qwen2_old.py:146-156).
 * scale_numel: 1 (per-tensor) or num_kv_heads (per-head); scales are device f32.
 */
B200ATTN_API int b200attn_reshape_and_cache(
    const void* key, const void* value,
    int64_t key_stride_tok, int64_t key_stride_head,
    int64_t value_stride_tok, int64_t value_stride_head,
    void* k_cache, void* v_cache,
    int64_t cache_stride_block, int64_t cache_stride_slot, int64_t cache_stride_head,
    const int64_t* slot_mapping, int32_t num_tokens,
    int32_t num_kv_heads, int32_t head_size, int32_t block_size, int64_t num_blocks,
    int32_t kv_dtype,
    const float* k_scale, const float* v_scale, int32_t scale_numel,
    void* stream);

/*
 * Fused rotary embedding + KV-cache write. Replaces the two launches the reference's attention
 * layers issue before every attention call — `q, k = self.rotary_emb(positions, q, k)`
 * (vllm_omni/model_executor/models/qwen2_5_omni/qwen2_old.py:151-153; vLLM
 * model_executor/layers/rotary_embedding/{base,mrope}.py) and the cache write above — behind vLLM's hook
 * AttentionImpl.do_rope_and_kv_cache_update (vllm/v1/attention/backend.py:815-840).
 *
 * For every token t < num_tokens, q head and k head: the first rot_dim elements are rotated IN PLACE
 * in bf16 arithmetic (every product and sum rounded to bf16, as torch computes on bf16 tensors):
 *     o1 = x1 * cos - x2 * sin,  o2 = x2 * cos + x1 * sin
 *     NeoX (is_neox != 0): (x1, x2) = (x[p], x[p + rot_dim/2]); GPT-J: (x[2p], x[2p+1]); p < rot_dim/2
 *     cos = cos_sin_cache[pos, p], sin = cos_sin_cache[pos, rot_dim/2 + p]   (bf16 rows, stride given)
 * positions: i64 [num_tokens] (pos_rows = 1) or [3, num_tokens] (pos_rows = 3, M-RoPE: pair p takes its
 * position from row r(p): chunked — r = section index of p in mrope_section (host i32[3]); interleaved —
 * r = 1 if p % 3 == 1 and p < 3 * section[1], r = 2 if p % 3 == 2 and p < 3 * section[2], else 0).
 * Then, for t < num_slot_tokens with slot_mapping[t] >= 0, the rotated key row and the value row are
 * written to the cache exactly as b200attn_reshape_and_cache would (bf16 bit copy / e4m3 of x / scale).
 */
B200ATTN_API int b200attn_rope_and_cache(
    void* query, int64_t q_stride_tok, int32_t num_q_heads,
    void* key, const void* value,
    int64_t key_stride_tok, int64_t key_stride_head,
    int64_t value_stride_tok, int64_t value_stride_head,
    const int64_t* positions, int32_t pos_rows, int64_t pos_stride_row,
    const void* cos_sin_cache, int64_t cos_sin_stride, int64_t max_position,
    int32_t rot_dim, int32_t is_neox,
    const int32_t* mrope_section /* host, 3 ints, or NULL */, int32_t mrope_interleaved,
    void* k_cache, void* v_cache,
    int64_t cache_stride_block, int64_t cache_stride_slot, int64_t cache_stride_head,
    const int64_t* slot_mapping, int32_t num_tokens, int32_t num_slot_tokens,
    int32_t num_kv_heads, int32_t head_size, int32_t block_size, int64_t num_blocks,
    int32_t kv_dtype,
    const float* k_scale, const float* v_scale, int32_t scale_numel,
    void* stream);

/*
 * Split-K heuristic used when num_splits == 0 (exported so the host shim can size the
 * workspace and so tests can pin it). Pure host arithmetic.
 */
B200ATTN_API int32_t b200attn_decode_num_splits(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                                   int32_t max_seq_len, int32_t num_sms);

/* Bytes of split-K workspace for (max_batch, num_q_heads, head_size, max_splits). */
B200ATTN_API size_t b200attn_decode_workspace_bytes(int32_t max_batch, int32_t num_q_heads,
                                       int32_t head_size, int32_t max_splits);

/*
 * One-time preparation of a decode workspace: zeroes its control area (the first 16 KiB:
 * stream-K claim / arrival counters). Call once after allocating the workspace (or allocate it
 * zero-filled); the library leaves the control area zero after every launch, so graphs can be
 * replayed without re-initialising. A workspace must be at least 16 KiB.
 */
B200ATTN_API int b200attn_workspace_init(void* workspace, size_t workspace_bytes, void* stream);

/*
 * Stream-K plan of one decode step (what FlashAttention's get_scheduler_metadata / FlashInfer's plan()
 * do for the stock backends: vllm/v1/attention/backends/flash_attn.py:388-575): tiles prefix per
 * sequence, the start position of every warp's chunk of the step's KV-tile stream and the block ids
 * it starts with, written into `workspace`. One small launch per STEP; every layer's
 * b200attn_paged_decode of that step then passes B200ATTN_DECODE_PLANNED with the same seq_lens,
 * block_table, batch, head counts, block_size and workspace. Without the flag b200attn_paged_decode
 * runs this launch itself before every decode kernel.
 */
B200ATTN_API int b200attn_decode_plan(
    const int32_t* seq_lens, int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
    const int32_t* block_table, int64_t block_table_stride, int32_t block_size,
    void* workspace, size_t workspace_bytes, void* stream);

/*
 * Paged decode attention (query length 1 per sequence). Replaces the decode branch of
 * AttentionImpl.forward in the reference's stock backends
 * (vllm/v1/attention/backends/flashinfer.py:1694-1822 — TRT-LLM-gen decode, and
 *  vllm/v1/attention/backends/flash_attn.py:758-819 — flash_attn_varlen_func with
 *  block_table/seqused_k) plus their split-KV combine.
 *
 * For each sequence b < batch and query head h:
 *     kvh = h / (num_q_heads / num_kv_heads)
 *     K, V = cache[block_table[b, 0:ceil(S_b/bs)]] flattened, first S_b = seq_lens[b] rows
 *     out[b, h, :] = softmax(softmax_scale * q[b,h,:] . K[:, kvh, :]^T) . V[:, kvh, :]
 *     lse[b, h]    = log(sum_j exp(softmax_scale * q.k_j))       (natural log; optional)
 * fp8 caches are dequantised as float(e4m3) * scale[kvh or 0].
 * S_b == 0 => out row = 0, lse = -inf.
 *
 * Schedules: with num_splits == 0 and variant AUTO (what the backend uses) the KV tiles of
 * the whole step are cut into equal contiguous chunks, one per warp of a persistent grid
 * ("stream-K", B200ATTN_VARIANT_SK2: no split count, balanced for ragged batches, any batch;
 * needs block_size = 16 * 2^k and the workspace of b200attn_decode_workspace_bytes, otherwise the
 * fixed-split schedule is used). Two launches: the decode kernel (online softmax per chunk, fp32
 * partial outputs + LSE into `workspace`; items that lie inside one chunk are written to `out`
 * directly) and the separate LSE-reduce kernel that merges the partials of the items a chunk
 * boundary cut. With num_splits >= 1 (or an explicit fixed-split variant) every sequence is cut
 * into that many equal KV ranges instead.
 *
 * Determinism: the stream-K chunk sizes follow per-SM streaming rates that the library measures
 * over the first launches on a workspace (8 plan launches by default, B200ATTN_SK2_CALIB=n; 0 =
 * equal shares) and then freezes: from then on the output is a deterministic function of the
 * inputs; before that, and between different workspaces, results can differ in the last bf16 bit
 * (a different but equally valid summation order).
 *
 * q:   bf16 [batch, num_q_heads, head_size], token stride given, head stride = head_size
 * out: bf16 [batch, num_q_heads, head_size], token stride given
 * block_table: i32 [batch, >=ceil(max_seq_len/bs)], row stride given
 * seq_lens: i32 [batch] (device; never read on the host)
 * max_seq_len: host upper bound used for the split heuristic only (longer sequences
 *              are still handled correctly: the last split extends to S_b)
 * num_splits: 0 = automatic (stream-K, or the split heuristic for the fixed-split variants),
 *             else forced (1 = no split, no reduce launch)
 * variant: B200ATTN_VARIANT_*
 * flags: B200ATTN_DECODE_* (0 = self-contained call, no assumption about what runs before it)
 */
B200ATTN_API int b200attn_paged_decode(
    const void* q, int64_t q_stride_tok,
    const void* k_cache, const void* v_cache,
    int64_t cache_stride_block, int64_t cache_stride_slot, int64_t cache_stride_head,
    int64_t num_blocks,
    void* out, int64_t out_stride_tok,
    float* lse,
    const int32_t* block_table, int64_t block_table_stride,
    const int32_t* seq_lens, int32_t batch, int32_t max_seq_len,
    int32_t num_q_heads, int32_t num_kv_heads, int32_t head_size, int32_t block_size,
    float softmax_scale, int32_t kv_dtype,
    const float* k_scale, const float* v_scale, int32_t scale_numel,
    int32_t num_splits, int32_t variant, int32_t flags,
    void* workspace, size_t workspace_bytes, void* stream);

/*
 * b200attn_paged_decode for callers that cannot allocate device memory (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc): no workspace argument, no cudaMalloc / cudaFree /
 * allocating library call inside. NO COUNTERPART IN THE
 * REFERENCE.
 *
 * Same arguments and result as b200attn_paged_decode with variant = B200ATTN_VARIANT_STREAM (the fixed-split
 * schedule: persistent decode kernel + LSE-reduce launch; bit-identical to that call for the same num_splits),
 * bf16 caches only. The split-KV partial rows go to a file-scope `static __device__` array of
 * B200ATTN_STATIC_SCRATCH_BYTES (8 MiB) inside the library: every GPU that loads the library holds those
 * 8 MiB until the process exits.
 *
 * Limit: batch * num_q_heads * num_splits <= B200ATTN_STATIC_MAX_ROWS partial rows (num_splits == 0 picks
 * b200attn_decode_num_splits' value; num_splits == 1 writes `out` directly and has no limit). Anything beyond
 * is rejected with -B200ATTN_EWORKSPACE and a message that names the limit; fp8 caches with
 * -B200ATTN_EUNSUPPORTED.
 *
 * The array is shared by every call on the device and keeps the previous call's contents. Nothing in it has
 * to start at zero (a call reads only the rows it wrote itself), but two calls that overlap on the GPU —
 * different streams, or different host threads on different streams — RACE on it: serialise them (one stream,
 * or events) or use b200attn_paged_decode with a workspace each.
 */
#define B200ATTN_STATIC_SCRATCH_BYTES 8388608
#define B200ATTN_STATIC_MAX_ROWS 16000
B200ATTN_API size_t b200attn_static_scratch_bytes(void);
B200ATTN_API int b200attn_paged_decode_static(
    const void* q, int64_t q_stride_tok,
    const void* k_cache, const void* v_cache,
    int64_t cache_stride_block, int64_t cache_stride_slot, int64_t cache_stride_head,
    int64_t num_blocks,
    void* out, int64_t out_stride_tok,
    float* lse,
    const int32_t* block_table, int64_t block_table_stride,
    const int32_t* seq_lens, int32_t batch, int32_t max_seq_len,
    int32_t num_q_heads, int32_t num_kv_heads, int32_t head_size, int32_t block_size,
    float softmax_scale, int32_t kv_dtype,
    int32_t num_splits, void* stream);

/*
 * The two launches of b200attn_paged_decode, exported separately for profiling sweeps.
 * `num_splits` must be >= 1 here unless variant is B200ATTN_VARIANT_STREAMK / _SK2 (which ignore it);
 * AUTO means B200ATTN_VARIANT_STREAM. partial writes workspace (and `out` rows finished by a
 * single KV range); reduce merges the rows that were cut. Both calls must be given the same
 * batch, head counts, num_splits, variant and workspace.
 */
B200ATTN_API int b200attn_paged_decode_partial(
    const void* q, int64_t q_stride_tok,
    const void* k_cache, const void* v_cache,
    int64_t cache_stride_block, int64_t cache_stride_slot, int64_t cache_stride_head,
    int64_t num_blocks,
    void* out, int64_t out_stride_tok,
    float* lse,
    const int32_t* block_table, int64_t block_table_stride,
    const int32_t* seq_lens, int32_t batch, int32_t max_seq_len,
    int32_t num_q_heads, int32_t num_kv_heads, int32_t head_size, int32_t block_size,
    float softmax_scale, int32_t kv_dtype,
    const float* k_scale, const float* v_scale, int32_t scale_numel,
    int32_t num_splits, int32_t variant, int32_t flags,
    void* workspace, size_t workspace_bytes, void* stream);

B200ATTN_API int b200attn_lse_reduce(
    void* out, int64_t out_stride_tok, float* lse,
    const int32_t* seq_lens, int32_t batch, int32_t max_seq_len,
    int32_t num_q_heads, int32_t num_kv_heads, int32_t head_size,
    int32_t num_splits, int32_t variant,
    const void* workspace, size_t workspace_bytes, void* stream);

/*
 * Causal varlen attention over the paged cache for query length >= 1 (prefill, chunked
 * prefill / extend, mixed batches). Correctness path for the non-decode tokens a
 * registered backend also receives (vllm/v1/attention/backends/flash_attn.py:758-819 with
 * max_query_len > 1). The new tokens' K/V must already be in the cache.
 * bf16 caches run on the tcgen05 / TMEM kernel; fp8 caches on the mma.sync kernel here, and on the tcgen05 kernel
 * through b200attn_paged_prefill_fp8_staged below.
 *
 * Sequence b owns query rows [query_start_loc[b], query_start_loc[b+1]); with
 * q_len = their count and S_b = seq_lens[b], query i (0-based) attends to cache
 * positions j <= S_b - q_len + i.
 */
B200ATTN_API int b200attn_paged_prefill(
    const void* q, int64_t q_stride_tok,
    const void* k_cache, const void* v_cache,
    int64_t cache_stride_block, int64_t cache_stride_slot, int64_t cache_stride_head,
    int64_t num_blocks,
    void* out, int64_t out_stride_tok,
    const int32_t* block_table, int64_t block_table_stride,
    const int32_t* seq_lens, const int32_t* query_start_loc,
    int32_t batch, int32_t num_tokens, int32_t max_query_len,
    int32_t num_q_heads, int32_t num_kv_heads, int32_t head_size, int32_t block_size,
    float softmax_scale, int32_t kv_dtype,
    const float* k_scale, const float* v_scale, int32_t scale_numel,
    void* stream);

/*
 * b200attn_paged_prefill for FP8 (e4m3) caches on the tcgen05 / TMEM kernel. Same computation, same reference
 * call site (vllm/v1/attention/backends/flash_attn.py:758-819 with an fp8 kv_cache_dtype; FlashInfer's fp8
 * prefill branch, vllm/v1/attention/backends/flashinfer.py:1514-1688). Two steps on `stream`:
 *   1. the e4m3 pages that the `batch` sequences own are rewritten as bf16 pages in `staging` (the e4m3 VALUES,
 *      which bf16 holds exactly; slots at or past seq_lens[b] as zeros) together with their own block table;
 *   2. the tcgen05 prefill kernel runs over the staged pages with k_scale folded into the softmax scale and
 *      v_scale into the final normalisation — the same arithmetic as a dequantised cache, without rounding the
 *      dequantised values.
 * The staging costs 5 bytes of HBM traffic per cached element per layer, 1-2 % of a prefill step; without a staging
 * buffer b200attn_paged_prefill serves fp8 caches on its mma.sync kernel.
 *
 * max_seq_len: host upper bound of seq_lens[0..batch) (REQUIRED here: it sizes the staged block table).
 * total_kv_tokens: host upper bound of sum(seq_lens), or 0 if unknown (then batch * max_seq_len is provisioned).
 * staging: device buffer of b200attn_prefill_staging_bytes(...) for the same arguments, 1 KiB aligned, contents
 *          undefined on entry and exit; one buffer can serve every layer of a step in turn (stream order).
 */
B200ATTN_API size_t b200attn_prefill_staging_bytes(int32_t batch, int32_t max_seq_len, int64_t total_kv_tokens,
                                                   int32_t num_kv_heads, int32_t head_size, int32_t block_size);
B200ATTN_API int b200attn_paged_prefill_fp8_staged(
    const void* q, int64_t q_stride_tok,
    const void* k_cache, const void* v_cache,
    int64_t cache_stride_block, int64_t cache_stride_slot, int64_t cache_stride_head,
    int64_t num_blocks,
    void* out, int64_t out_stride_tok,
    const int32_t* block_table, int64_t block_table_stride,
    const int32_t* seq_lens, const int32_t* query_start_loc,
    int32_t batch, int32_t num_tokens, int32_t max_query_len,
    int32_t max_seq_len, int64_t total_kv_tokens,
    int32_t num_q_heads, int32_t num_kv_heads, int32_t head_size, int32_t block_size,
    float softmax_scale,
    const float* k_scale, const float* v_scale, int32_t scale_numel,
    void* staging, size_t staging_bytes, void* stream);

/*
 * Gather of one request's cached K or V rows into a dense [seq_len, num_kv_heads,
 * head_size] tensor (same dtype as the cache, no dequantisation). Device-side twin of the
 * reference's per-layer Python indexing in
 * vllm_omni/distributed/omni_connectors/kv_transfer_manager.py:829-846
 * (key_blocks[block_ids].flatten(0,1)[:seq_len]).
 * block_ids: i32 [num_block_ids] (device).
 */
B200ATTN_API int b200attn_gather_kv(
    const void* cache, int64_t cache_stride_block, int64_t cache_stride_slot,
    int64_t cache_stride_head, int64_t num_blocks,
    const int32_t* block_ids, int32_t num_block_ids, int32_t seq_len,
    int32_t num_kv_heads, int32_t head_size, int32_t block_size, int32_t kv_dtype,
    void* dst, void* stream);

/* ------------------------------------------------------------------------------------------
 * Peer-memory all-reduce (sum over ranks, bf16 in / fp32 accumulate in rank order / bf16 out).
 *
 * Replaces the collective behind the o_proj row-parallel sum that follows attention:
 * vllm/distributed/communication_op.py:12 tensor_model_parallel_all_reduce (reached from
 * RowParallelLinear.forward, called by Qwen2Attention.forward
 * vllm_omni/model_executor/models/qwen2_5_omni/qwen2_old.py:155 `output, _ = self.o_proj(attn_output)`).
 * One process per GPU; every rank owns one peer-mapped allocation (same size on every rank) and
 * passes HOST arrays of `world` device pointers — entry r = rank r's region as mapped into THIS
 * process. The launch is asynchronous on `stream`, CUDA-graph capturable, and touches no host state.
 *
 *   peer_flags[r]   B200ATTN_PEER_FLAG_BYTES of uint32 flag words, zeroed ONCE before the first
 *                   launch (then a cross-rank barrier); never reset afterwards.
 *   peer_in[r]      the bf16 input of rank r (16-byte aligned).
 *   peer_scratch[r] >= ceil(numel/8/world + 1) * 16 bytes, used by TWOSHOT (may be NULL otherwise).
 *   out             local result. ONESHOT/TWOSHOT: any device pointer that no peer reads (it may
 *                   not alias peer_in). MULTIMEM: this rank's unicast view of the region `mc_out`
 *                   multicasts to.
 *   peer_push[r]    rank r's receive region of the PUSH / PUSH2 schedules (or NULL array: unavailable):
 *                   3 x world x push_slot_bytes, filled ONCE with the bf16 pattern 0x8000 (-0.0, the
 *                   "not arrived" sentinel) before the first launch. PUSH reads its input only from
 *                   peer_in[rank], which may then be ANY local device pointer (no peer mapping needed).
 *   mc_in / mc_out  multicast addresses of the input / output regions (MULTIMEM) or NULL.
 *   numel           multiple of 8.  algo: AUTO picks by size (b200attn_peer_allreduce_algo).
 *   timeout_ns      a flag wait longer than this records {1, block, peer, epoch} in the 4 words after the
 *                   flag slots and traps the kernel (0 = 10 s). With B200ATTN_PEER_SOFT_TIMEOUT=1 in the
 *                   environment (tests) the wait is abandoned instead and the result is undefined.
 * ONESHOT, TWOSHOT and PUSH are bit-identical on every rank and equal to the sequential fp32 sum in rank
 * order rounded once to bf16; MULTIMEM leaves the order to the switch (fp32 accumulate).
 */
#define B200ATTN_PEER_FLAG_BYTES 8192
#define B200ATTN_PEER_MAX_RANKS 8
#define B200ATTN_AR_AUTO 0
#define B200ATTN_AR_ONESHOT 1
#define B200ATTN_AR_TWOSHOT 2
#define B200ATTN_AR_MULTIMEM 3
#define B200ATTN_AR_PUSH 4
#define B200ATTN_AR_PUSH2 5
/* pure host arithmetic: the schedule AUTO resolves to for `bytes` of input on `world` ranks */
B200ATTN_API int32_t b200attn_peer_allreduce_algo(int64_t bytes, int32_t world, int32_t has_multicast,
                                                  int64_t push_slot_bytes);
B200ATTN_API int b200attn_peer_allreduce(
    const void* const* peer_in, void* const* peer_scratch, void* const* peer_flags,
    void* const* peer_push, int64_t push_slot_bytes,
    void* out, const void* mc_in, void* mc_out,
    int64_t numel, int32_t world, int32_t rank, int32_t algo, int32_t max_blocks,
    uint64_t timeout_ns, void* stream);

/*
 * Row exchange inside a KV-head replica group (tp > KV heads, vllm_omni_b200.tp.ReplicaSplit) through
 * peer memory — replaces the two NCCL all_to_all calls around the decode. The reference rule
 * (qwen2_old.py:98-106) replicates a KV head on R ranks; here each of them serves ALL query heads of the
 * group for 1/R of the sequences, so q rows are pulled from the peers before the decode and output rows
 * are pushed back after it.
 *   peer_rows[g]  rank g's [batch, row_bytes] buffer (its query heads of every sequence: q, or out)
 *   local_rows    [rows_per_rank, group_size, row_bytes] staging the decode kernel reads / writes
 *   scatter = 0   gather : local[i][g] = peer_rows[g][first_row + i]  (one flag barrier, then pulls)
 *   scatter = 1   scatter: peer_rows[g][first_row + i] = local[i][g]  (pushes, then one fenced barrier:
 *                 when the launch retires on a rank, every peer's rows for it have landed)
 * i < num_rows <= rows_per_rank; rows_per_rank must be the same on every rank of the group.
 * gather(l) and scatter(l) must be issued pairwise by every rank of the group (the hazards between
 * layers are ordered by these two barriers). Flags: a B200ATTN_PEER_FLAG_BYTES region per rank that
 * no other collective uses.
 */
B200ATTN_API int b200attn_peer_exchange(
    void* const* peer_rows, void* const* peer_flags, void* local_rows,
    int32_t group_size, int32_t group_rank, int32_t first_row, int32_t num_rows, int32_t rows_per_rank,
    int64_t row_bytes, int32_t scatter, uint64_t timeout_ns, void* stream);

/*
 * Peer-mappable allocations through CUDA IPC for hosts without torch symmetric memory:
 * alloc = cudaMalloc + zero fill + cudaIpcGetMemHandle (64-byte handle to ship to the peers by any
 * host channel), open = cudaIpcOpenMemHandle with lazy peer access, close / free undo them.
 */
B200ATTN_API int b200attn_peer_alloc(size_t bytes, void** ptr, void* handle64);
B200ATTN_API int b200attn_peer_open(const void* handle64, void** ptr);
B200ATTN_API int b200attn_peer_close(void* ptr);
B200ATTN_API int b200attn_peer_free(void* ptr);

/*
 * Debug aid for profiling the stream-K schedule: when set (thread-local, device pointer to
 * 3 x 8 bytes per warp of the persistent grid, or NULL to disable), every warp of the next
 * stream-K decode launches records its {entry, static-chunk end, end} %globaltimer stamps.
 */
B200ATTN_API void b200attn_debug_set_trace(void* device_buffer);

/*
 * Debug / test aid: force the team size (warps that share an item: 1, 2, 3, 4, 6 or 12) the next
 * cooperative stream-K plans of this thread use; 0 restores the plan kernel's own rule (from the
 * step's average item length). Results do not depend on it beyond fp32 summation order.
 */
B200ATTN_API void b200attn_debug_set_team(int32_t warps);

/* TMA tensor maps this thread has encoded so far. The library caches one map per (cache pointer, strides, shape) in a
 * 1024-entry 8-way table; in steady state (the layers' caches do not move) this number stops growing — a caller that
 * sees it grow per step is re-encoding maps on the host at every launch. */
B200ATTN_API uint64_t b200attn_debug_map_encodes(void);

/* Byte offset, inside a decode workspace, of the plan b200attn_decode_plan writes for (batch, num_q_heads)
 * on the current device (int32 words: header 16 {positions, CTAs G, team size T, teams, batch, head groups,
 * 0, magic, teams per CTA} | team starts [12 G + 1, padded to 4] | team descriptors [12 G][4] | team info
 * [12 G][4] | warp descriptors [12 G][4] | tiles prefix [batch + 1]); tests read the plan back through it. */
B200ATTN_API size_t b200attn_debug_plan_offset(int32_t batch, int32_t num_q_heads);

/*
 * The stream-K schedule for `total_tiles` KV tiles on `num_warps` warps (pure host arithmetic,
 * the same function the kernels evaluate): out5 = {static region S, static tiles per warp,
 * tiles per dynamic chunk, dynamic chunks, dynamic enabled}. Exported so tests can pin it.
 */
B200ATTN_API void b200attn_debug_sk_schedule(int64_t total_tiles, int32_t num_warps, int32_t dyn_shift,
                                             int64_t* out5);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
