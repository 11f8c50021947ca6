/*
 * b200_paged_attn.h -- C ABI of the B200-native paged-decode hot path.
 *
 * This is the drop-in boundary: plain pointers, sizes and enums, no torch types.
 * Each entry point replaces one launcher of the reference (paths relative to the
 * reference repository root):
 *
 *   b200pa_paged_attention_v1       csrc/attention/paged_attention_v1.cu:160-182
 *                                   (launcher :46-125; schema csrc/torch_bindings.cpp:52-62)
 *   b200pa_paged_attention_v2       csrc/attention/paged_attention_v2.cu:167-192
 *                                   (launcher :47-126; schema csrc/torch_bindings.cpp:66-76)
 *   b200pa_reshape_and_cache        csrc/cache_kernels.cu:416-442
 *                                   (schema csrc/torch_bindings.cpp:667-673)
 *   b200pa_reshape_and_cache_flash  csrc/cache_kernels.cu:459-497
 *                                   (schema csrc/torch_bindings.cpp:676-682)
 *   b200pa_convert_fp8              csrc/cache_kernels.cu:573-623
 *                                   (schema csrc/torch_bindings.cpp:696-699)
 *   b200pa_get_max_shared_memory_per_block_device_attribute
 *                                   csrc/cuda_utils_kernels.cu (schema
 *                                   csrc/torch_bindings.cpp:718-730)
 *   b200pa_swap_blocks              csrc/cache_kernels.cu:27-70   (schema torch_bindings.cpp:652-654)
 *   b200pa_copy_blocks              csrc/cache_kernels.cu:125-172 (kernel :74-101; schema :657-660)
 *   b200pa_gather_and_maybe_dequant_cache
 *                                   csrc/cache_kernels.cu:734-782 (kernel :628-712; schema :703-710)
 *   b200pa_cp_gather_cache          csrc/cache_kernels.cu:865-926 (kernel :785-849; schema :712-715)
 *   b200pa_merge_attn_states        csrc/attention/merge_attn_states.cu:172-180
 *                                   (launcher :131-170; schema csrc/torch_bindings.cpp:80-87)
 *   b200pa_correct_attn_out         vllm/attention/ops/common.py:86-116 (Triton kernel :9-68)
 *   b200pa_prefill_attention        flash_attn_varlen_func with block_table as called by
 *                                   vllm/v1/attention/backends/cpu_attn.py:584-612; Triton
 *                                   context_attention_fwd (vllm/attention/ops/paged_attn.py:200-242)
 *   b200pa_paged_attention_lse      no reference kernel: decode + log-sum-exp, what
 *                                   AttentionImpl.can_return_lse_for_decode
 *                                   (vllm/attention/backends/abstract.py:115-138) asks of a
 *                                   backend for decode context parallelism
 *   b200pa_paged_attention_v1_builtin_scratch
 *                                   no counterpart in the reference: paged_attention_v1 with the
 *                                   scratch inside the library, for callers that cannot allocate
 *
 * Conventions
 *   - All data pointers are DEVICE pointers on the current CUDA device; strides
 *     are in ELEMENTS of the pointed-to type, exactly the values the reference
 *     launchers read from the tensors (query.stride(0), key_cache.stride(0), ...).
 *   - k_scale / v_scale are DEVICE pointers to one fp32 (read in-kernel; no host
 *     sync, CUDA-graph capturable), as in the reference kernels.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - Every launcher is asynchronous, allocates nothing, keeps no state, and
 *     mutates only the buffers documented as outputs (and the caller-provided
 *     `workspace` of the attention entry points).
 *   - Return 0 on success.  Non-zero = the configuration is outside the supported
 *     domain or the launch failed; b200pa_last_error() then returns a message
 *     (thread-local, valid until the next call on the same thread).  The torch
 *     shim turns that into TORCH_CHECK(false, msg) -> Python RuntimeError, which
 *     is the reference's error convention (paged_attention_v1.cu:122,156).
 *   - There is NO CPU fallback: without a CUDA device every compute entry point
 *     fails with B200PA_ERR_CUDA.
 */
#ifndef B200_PAGED_ATTN_H_
#define B200_PAGED_ATTN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200PA_ABI_VERSION 8

/* scalar_t of query / out / key / value (reference: VLLM_DISPATCH_FLOATING_TYPES) */
enum { B200PA_F32 = 0, B200PA_F16 = 1, B200PA_BF16 = 2 };
/* vllm::Fp8KVCacheDataType (csrc/attention/dtype_fp8.cuh:15-19) */
enum { B200PA_KV_AUTO = 0, B200PA_KV_FP8_E4M3 = 1, B200PA_KV_FP8_E5M2 = 2 };

enum {
  B200PA_OK = 0,
  B200PA_ERR_UNSUPPORTED = 1, /* head size / block size / dtype outside the domain */
  B200PA_ERR_INVALID = 2,     /* inconsistent arguments */
  B200PA_ERR_CUDA = 3         /* CUDA runtime error (no device, launch failure) */
};

int b200pa_abi_version(void);
const char* b200pa_last_error(void);

/* "auto" | "fp8" | "fp8_e4m3" | "fp8_e5m2" -> B200PA_KV_*; -1 if unknown
 * (DISPATCH_BY_KV_CACHE_DTYPE, quant_utils.cuh:545-582). */
int b200pa_kv_dtype_from_string(const char* kv_cache_dtype);

/* (The process-wide test hooks -- kernel-family override, tuning variant, in-kernel trace --
 * are NOT part of this ABI: include/b200_paged_attn_testing.h, compiled in only with
 * -DB200PA_TESTING=1.) */
/* Number of kernels launched by this library so far in this process. */
uint64_t b200pa_launch_count(void);

/*
 * Bytes of device scratch the attention entry points want in `workspace` for head size
 * `head_size` on the current device (a function of the SM count only, about 20 MB on a B200;
 * independent of batch and context).  Contract (ABI v7):
 *   - its first B200PA_WORKSPACE_HEADER_BYTES must be ZERO before the first call that uses it
 *     (cudaMemsetAsync once, at allocation); every call leaves them zero again -- they hold the
 *     ticket counters with which the decode kernel's CTAs elect the one that merges a sequence
 *     cut by CTA boundaries (the reference needs a second kernel for that:
 *     paged_attention_v2_reduce_kernel, attention_kernels.cuh:560-664);
 *   - the rest needs no initialisation;
 *   - it may be shared by calls of any shape on ONE stream (calls that may run concurrently
 *     need one workspace each) and must be 16-byte aligned.
 * The torch shim keeps one zero-initialised workspace per (device, stream).  workspace == NULL is
 * allowed: the launchers then use the generic SIMT kernel, which needs none (and is several
 * times slower).
 */
#define B200PA_WORKSPACE_HEADER_BYTES 4096
int64_t b200pa_paged_attention_workspace_bytes(int head_size);

/*
 * out[num_seqs, num_heads, head_size]                      (written)
 * query[num_seqs, num_heads, head_size], row stride q_stride
 * key_cache[num_blocks, num_kv_heads, head_size/x, block_size, x]
 * value_cache[num_blocks, num_kv_heads, head_size, block_size]
 *   x = 16 / sizeof(cache element); strides kv_block_stride, kv_head_stride
 * block_tables[num_seqs, max_num_blocks_per_seq] int32, seq_lens[num_seqs] int32
 * alibi_slopes[num_heads] fp32 or NULL
 */
int b200pa_paged_attention_v1(
    void* out, const void* query, const void* key_cache, const void* value_cache,
    int num_seqs, int num_heads, int num_kv_heads, int head_size, int block_size,
    float scale, const int32_t* block_tables, int max_num_blocks_per_seq,
    const int32_t* seq_lens, int max_seq_len, const float* alibi_slopes,
    int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale,
    int tp_rank, int blocksparse_local_blocks, int blocksparse_vert_stride,
    int blocksparse_block_size, int blocksparse_head_sliding_step,
    void* workspace, int64_t workspace_bytes, void* stream);

/*
 * b200pa_paged_attention_v1 for a caller that cannot give it a workspace.  (No counterpart in the
 * reference: its kernels take all scratch from torch's allocator, which such a caller does not
 * have.)
 *
 * Who this is for.  The entry points above leave allocation to the caller: you size the workspace
 * with b200pa_paged_attention_workspace_bytes, cudaMalloc it, zero its header, and pass it in.
 * Some hosts of this ABI cannot do that --
 *   - a plugin inside an engine that handed every free byte of HBM to its KV-cache pool at start-up
 *     (the reference sizes the pool from what is left after profiling), where a cudaMalloc made
 *     later may fail and all device memory is owned by the engine's arena;
 *   - an FFI host (cgo, JNI, ctypes) that only ever receives device pointers from another
 *     component and has no CUDA runtime binding to allocate with;
 *   - a decode loop that must not run into cudaMalloc's device-wide synchronisation.
 * For them the scratch is a static device array compiled into libb200pa.so.  It costs
 * B200PA_BUILTIN_SCRATCH_BYTES (24 MiB) of device memory on EVERY GPU the library's code is loaded
 * on, from load until the process exits, used or not.
 *
 * Arguments, domain, result and error codes: exactly b200pa_paged_attention_v1 without
 * `workspace` / `workspace_bytes`; the same kernels run, so the output equals that entry point's.
 * The call itself allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).
 *
 * Limits and rules:
 *   - The scratch the kernels need depends on head_size and the SM count only.  24 MiB covers
 *     head_size <= 128 on a B200 (148 SMs).  A larger requirement (head_size 192 / 256, or a part
 *     with more than 151 SMs) is refused with B200PA_ERR_UNSUPPORTED and a message that names both
 *     sizes; use b200pa_paged_attention_v1 with your own workspace there.
 *   - The FIRST call on each device looks the array's device address up (once per device, with that
 *     device current -- make the call with the device you will run on selected).  Make that first
 *     call during start-up, before any stream capture.
 *   - One array per GPU, shared by every call of this entry point from every thread, and it keeps
 *     whatever the last call left in it.  Each call clears the 4 KiB of ticket counters at its
 *     front on `stream` before launching.  Two calls that can overlap on the GPU (different
 *     streams without an event between them) RACE on it and produce wrong output: serialise them,
 *     or give the concurrent ones their own workspace through b200pa_paged_attention_v1.
 */
#define B200PA_BUILTIN_SCRATCH_BYTES (24 * 1024 * 1024)
int b200pa_paged_attention_v1_builtin_scratch(
    void* out, const void* query, const void* key_cache, const void* value_cache,
    int num_seqs, int num_heads, int num_kv_heads, int head_size, int block_size,
    float scale, const int32_t* block_tables, int max_num_blocks_per_seq,
    const int32_t* seq_lens, int max_seq_len, const float* alibi_slopes,
    int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale,
    int tp_rank, int blocksparse_local_blocks, int blocksparse_vert_stride,
    int blocksparse_block_size, int blocksparse_head_sliding_step, void* stream);

/*
 * Split-K variant.  Caller-provided scratch, reference shapes:
 *   exp_sums, max_logits [num_seqs, num_heads, max_num_partitions] fp32
 *   tmp_out              [num_seqs, num_heads, max_num_partitions, head_size] scalar_t
 * with max_num_partitions >= ceil(max_seq_len / 512)  (_PARTITION_SIZE,
 * vllm/attention/ops/paged_attn.py:21).  `out` is the contract.  The scratch is
 * working memory: the stream kernel splits the context by its own schedule and
 * keeps its partials in `workspace`, leaving the three buffers untouched; the
 * generic kernel leaves in every slot j < ceil(seq_len/512) of a (seq, head) the
 * reference's partial (normalised partial output, its exp-sum and max logit).
 */
int b200pa_paged_attention_v2(
    void* out, float* exp_sums, float* max_logits, void* tmp_out,
    const void* query, const void* key_cache, const void* value_cache,
    int num_seqs, int num_heads, int num_kv_heads, int head_size, int block_size,
    float scale, const int32_t* block_tables, int max_num_blocks_per_seq,
    const int32_t* seq_lens, int max_seq_len, int max_num_partitions,
    const float* alibi_slopes, int64_t q_stride, int64_t kv_block_stride,
    int64_t kv_head_stride, int scalar_type, int kv_dtype, const float* k_scale,
    const float* v_scale, int tp_rank, int blocksparse_local_blocks,
    int blocksparse_vert_stride, int blocksparse_block_size,
    int blocksparse_head_sliding_step, void* workspace, int64_t workspace_bytes,
    void* stream);

/*
 * paged_attention_v1 (dense, no block-sparse) that also returns, per (sequence, head), the
 * natural-log sum-exp of the scaled logits:
 *   lse[seq * lse_seq_stride + head * lse_head_stride] = log(sum_t exp(scale * q.k_t [+ alibi]))
 * (-inf for an empty sequence; strides in floats, so both [num_seqs, num_heads] and the
 * [num_heads, num_seqs] layout of merge_attn_states are expressible).
 */
int b200pa_paged_attention_lse(
    void* out, float* lse, int64_t lse_seq_stride, int64_t lse_head_stride,
    const void* query, const void* key_cache, const void* value_cache,
    int num_seqs, int num_heads, int num_kv_heads, int head_size, int block_size,
    float scale, const int32_t* block_tables, int max_num_blocks_per_seq,
    const int32_t* seq_lens, int max_seq_len, const float* alibi_slopes,
    int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale,
    void* workspace, int64_t workspace_bytes, void* stream);

/*
 * One decode step of one attention layer in ONE launch: the effect of
 *   b200pa_reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, num_tokens = num_seqs)
 * followed by b200pa_paged_attention_v1 (and the lse store of b200pa_paged_attention_lse when
 * lse != NULL), i.e. PagedAttention.write_to_paged_cache + forward_decode of the reference
 * (vllm/attention/ops/paged_attn.py:73-92, 94-198) for a pure decode batch.
 *   key, value [num_seqs, num_kv_heads, head_size] scalar_t, row strides key_stride /
 *   value_stride; slot_mapping[num_seqs] int64 (< 0: token neither stored nor patched in).
 * PRECONDITION (what a decode batch is): token b is position seq_lens[b] - 1 of sequence b, so
 * slot_mapping[b] == block_tables[b][(seq_lens[b]-1) / block_size] * block_size +
 * (seq_lens[b]-1) % block_size, and seq_lens[b] >= 1.  The kv cache must be dense.
 * Where the stream kernel serves the shape (and key/value rows are 16-byte aligned) the warp
 * that reaches the end of a (sequence, kv head) takes the new token from key/value, quantised
 * exactly as reshape_and_cache does, and stores it to the cache; otherwise the two kernels are
 * launched back to back.  The resulting cache bytes and attention output are the same as those
 * of the two separate calls.
 */
int b200pa_decode_step(
    void* out, float* lse, int64_t lse_seq_stride, int64_t lse_head_stride, const void* query,
    const void* key, const void* value, int64_t key_stride, int64_t value_stride,
    const int64_t* slot_mapping, void* key_cache, void* value_cache,
    int num_seqs, int num_heads, int num_kv_heads, int head_size, int block_size,
    float scale, const int32_t* block_tables, int max_num_blocks_per_seq,
    const int32_t* seq_lens, int max_seq_len, const float* alibi_slopes,
    int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale,
    void* workspace, int64_t workspace_bytes, void* stream);

/*
 * Decode attention over a TOKEN-MAJOR ("flash") cache -- the layout reshape_and_cache_flash
 * writes (csrc/cache_kernels.cu:280-353 of the reference) and every flash-attention style
 * backend and KV connector of the reference uses -- so that the decode kernel serves it without
 * a layout conversion:
 *   key_cache, value_cache: [num_blocks][block_size][num_kv_heads][head_size] by strides
 *   (elements): block_stride, page_stride (between two tokens of a block), head_stride;
 *   NHD: page_stride = num_kv_heads*head_size, head_stride = head_size;
 *   HND: head_stride = block_size*head_size, page_stride = head_size.
 * Same math, same output contract as b200pa_paged_attention_v1 / _lse (lse may be NULL).
 * key / value / slot_mapping all non-NULL: additionally the effect of
 * b200pa_reshape_and_cache_flash(key, value, ..., slot_mapping, num_tokens = num_seqs) BEFORE the
 * attention, in the same launch, under b200pa_decode_step's precondition (token b is position
 * seq_lens[b] - 1 of sequence b).  All three NULL: attention only.
 * Domain: head_size 32 / 64 / 80 / 96 / 112 / 128 / 192 / 256, block_size 16 / 32, f16 / bf16
 * query, auto / fp8_e4m3 / fp8_e5m2 cache, 16-byte aligned rows; anything else is
 * B200PA_ERR_UNSUPPORTED (there is no second kernel for this layout).  A workspace
 * (b200pa_paged_attention_workspace_bytes) is required.  NHD caches with an even number of kv
 * heads, head size 128 and 16-token blocks are read a whole block (all heads) at a time through
 * tensor-map TMA (csrc/paged_attention_nhd.cuh); everything else by rows.
 */
int b200pa_paged_attention_flash(
    void* out, float* lse, int64_t lse_seq_stride, int64_t lse_head_stride, const void* query,
    const void* key, const void* value, int64_t key_stride, int64_t value_stride,
    const int64_t* slot_mapping, void* key_cache, void* value_cache,
    int num_seqs, int num_heads, int num_kv_heads, int head_size, int block_size,
    float scale, const int32_t* block_tables, int max_num_blocks_per_seq,
    const int32_t* seq_lens, int max_seq_len, const float* alibi_slopes,
    int64_t q_stride, int64_t block_stride, int64_t page_stride, int64_t head_stride,
    int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale,
    void* workspace, int64_t workspace_bytes, void* stream);

/*
 * Causal attention of query CHUNKS over the paged cache -- prefill, chunked prefill and
 * multi-token (speculative) decode; the second kernel the reference's paged backends hand every
 * batch with more than one query token per sequence to: flash_attn_varlen_func(out, query,
 * key_cache, value_cache, cu_seqlens_q, cu_seqlens_kv, ..., causal, block_table, alibi_slopes) in
 * vllm/v1/attention/backends/cpu_attn.py:584-612, context_attention_fwd behind
 * PagedAttention.forward_prefix (vllm/attention/ops/paged_attn.py:200-242).
 *   query, out [num_tokens, num_heads, head_size] scalar_t (token strides q_stride / out_stride,
 *   heads dense); lse [num_tokens, num_heads] fp32 natural-log sum-exp or NULL;
 *   query_start_loc[num_seqs + 1] int32 (tokens of sequence b: [qsl[b], qsl[b+1]));
 *   seq_lens[num_seqs] int32 = context + new tokens; the K/V of the new tokens are already in the
 *   cache (write it first: b200pa_reshape_and_cache).  Query token i of sequence b sits at
 *   position ctx + i, ctx = seq_lens[b] - q_len, and attends cache positions [0, ctx + i].
 *   key_cache / value_cache: the x-packed layout of b200pa_paged_attention_v1.
 *   heads_per_tile: query heads of one kv head stacked per token in a 64-row tile (must divide
 *   num_heads / num_kv_heads); 0 = choose (the whole group when max_query_len < 64, else 1).
 * Domain: head_size 32 / 64 / 96 / 128 / 192 / 256, block_size 16 / 32, f16 / bf16, kv auto /
 * fp8_e4m3 / fp8_e5m2; 16-byte aligned query rows.  No workspace.
 */
int b200pa_prefill_attention(
    void* out, float* lse, const void* query, const void* key_cache, const void* value_cache,
    int num_seqs, int num_heads, int num_kv_heads, int head_size, int block_size, float scale,
    const int32_t* block_tables, int max_num_blocks_per_seq, const int32_t* query_start_loc,
    const int32_t* seq_lens, int max_query_len, const float* alibi_slopes, int64_t q_stride,
    int64_t out_stride, int64_t kv_block_stride, int64_t kv_head_stride, int scalar_type,
    int kv_dtype, const float* k_scale, const float* v_scale, int heads_per_tile, void* stream);

/*
 * output, prefix_output, suffix_output [num_tokens, num_heads, head_size] scalar_t, dense,
 * 16-byte aligned; prefix_lse, suffix_lse, output_lse (may be NULL) [num_heads, num_tokens] fp32.
 * head_size must be a multiple of 16 / sizeof(scalar_t).
 */
int b200pa_merge_attn_states(
    void* output, float* output_lse, const void* prefix_output, const float* prefix_lse,
    const void* suffix_output, const float* suffix_lse, int64_t num_tokens, int num_heads,
    int head_size, int scalar_type, void* stream);

/*
 * Decode-context-parallel correction, in place:
 *   lse_total = log sum_n exp(lses[n, b, h]);  out[b, h, :] *= exp(lses[rank, b, h] - lse_total)
 * out [num_seqs, num_heads, head_size] scalar_t with element strides out_seq_stride /
 * out_head_stride (innermost dense); lses [num_ranks, num_seqs, num_heads] fp32 by strides;
 * lse_out (written) is indexed with lses' seq / head strides.
 */
int b200pa_correct_attn_out(
    void* out, float* lse_out, const float* lses, int num_ranks, int rank, int num_seqs,
    int num_heads, int head_size, int64_t out_seq_stride, int64_t out_head_stride,
    int64_t lses_rank_stride, int64_t lses_seq_stride, int64_t lses_head_stride,
    int scalar_type, void* stream);

/*
 * Block management on the same caches (the reference's _C_cache_ops.swap_blocks / copy_blocks /
 * gather_and_maybe_dequant_cache / cp_gather_cache).
 */
enum { B200PA_COPY_D2D = 0, B200PA_COPY_D2H = 1, B200PA_COPY_H2D = 2 };

/*
 * For every pair (s, d) of block_mapping (HOST memory, [num_pairs][2] int64, as the reference
 * requires a CPU tensor): copy block_bytes bytes from src + s*block_bytes to dst + d*block_bytes,
 * asynchronously on `stream` (cudaMemcpyAsync; block_bytes = element_size * src.stride(0), padding
 * included).  Runs of consecutive (s+1, d+1) pairs are issued as one copy.
 */
int b200pa_swap_blocks(const void* src, void* dst, int64_t block_bytes,
                       const int64_t* block_mapping_host, int64_t num_pairs, int direction,
                       void* stream);

/*
 * key_cache_ptrs / value_cache_ptrs: HOST arrays of num_layers DEVICE pointers; block_mapping:
 * DEVICE [num_pairs][2] int64.  In every layer, block s of the key cache is copied onto block d of
 * the key cache, same for the value cache; block_bytes = numel of one block * element size (the
 * reference's numel_per_block, no padding).  No allocation and no host synchronisation (the
 * reference uploads the pointer arrays with a blocking copy).
 */
int b200pa_copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs, int num_layers,
                       const int64_t* block_mapping, int num_pairs, int64_t block_bytes,
                       void* stream);

/*
 * src_cache [num_blocks, block_size, entry...] cache_t by strides (elements), dst [tot_tokens,
 * entry_size] scalar_t (row stride dst_entry_stride), block_table [batch, *] int32 (row stride
 * block_table_stride), cu_seq_lens [batch + 1] int32, seq_starts [batch] int32 or NULL
 * (seq_starts[b] / block_size whole blocks are skipped at the front of row b), scale: device fp32
 * (read for fp8 caches only).  kv_dtype auto: cache_t == scalar_t, plain gather; fp8: bytes ->
 * scalar_t(float(half(fp8)) * scale).  tot_tokens (dst rows; only used to size the grid -- the
 * kernel reads the true total from cu_seq_lens[batch]).
 */
int b200pa_gather_and_maybe_dequant_cache(
    const void* src_cache, void* dst, const int32_t* block_table, const int32_t* cu_seq_lens,
    int batch_size, int block_size, int entry_size, int64_t block_table_stride,
    int64_t cache_block_stride, int64_t cache_entry_stride, int64_t dst_entry_stride,
    int scalar_type, int kv_dtype, const float* scale, const int32_t* seq_starts,
    int64_t tot_tokens, void* stream);

/*
 * Same gather without conversion for elements of element_bytes in {1, 2, 4}; seq_starts counts
 * TOKENS and need not be a multiple of block_size.
 */
int b200pa_cp_gather_cache(
    const void* src_cache, void* dst, const int32_t* block_table, const int32_t* cu_seq_lens,
    int batch_size, int block_size, int entry_size, int64_t block_table_stride,
    int64_t cache_block_stride, int64_t cache_entry_stride, int64_t dst_entry_stride,
    int element_bytes, const int32_t* seq_starts, int64_t tot_tokens, void* stream);

/*
 * key, value [num_tokens, num_heads, head_size] scalar_t, row strides key_stride /
 * value_stride; slot_mapping[num_tokens] int64 (slot < 0: token skipped);
 * caches in the x-packed layout above (dense, as the reference indexes them).
 */
int b200pa_reshape_and_cache(
    const void* key, const void* value, void* key_cache, void* value_cache,
    const int64_t* slot_mapping, int num_tokens, int64_t key_stride,
    int64_t value_stride, int num_heads, int head_size, int block_size, int x,
    int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale,
    void* stream);

/*
 * Flash layout, addressed by strides only: NHD [num_blocks, block_size, num_heads,
 * head_size] (head_stride == head_size) or HND [num_blocks, num_heads, block_size,
 * head_size].  block_stride is shared by key_cache and value_cache.
 */
int b200pa_reshape_and_cache_flash(
    const void* key, const void* value, void* key_cache, void* value_cache,
    const int64_t* slot_mapping, int num_tokens, int64_t key_stride,
    int64_t value_stride, int64_t block_stride, int64_t page_stride,
    int64_t head_stride, int num_heads, int head_size, int block_size,
    int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale,
    void* stream);

/*
 * Whole-tensor fp8 <-> scalar_t conversion with a HOST scale (test utility).
 * to_fp8 != 0: src scalar_t -> dst uint8 codes; else src uint8 -> dst scalar_t.
 * kv_dtype must be B200PA_KV_FP8_E4M3 or B200PA_KV_FP8_E5M2.
 */
int b200pa_convert_fp8(void* dst, const void* src, int64_t num_elements,
                       float scale, int scalar_type, int kv_dtype, int to_fp8,
                       void* stream);

/* cudaDevAttrMaxSharedMemoryPerBlockOptin of `device_id`. */
int b200pa_get_max_shared_memory_per_block_device_attribute(int device_id,
                                                            int64_t* value);

#ifdef __cplusplus
}
#endif
#endif /* B200_PAGED_ATTN_H_ */
