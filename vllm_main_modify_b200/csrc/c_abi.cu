// extern "C" entry points of libb200pa.so (see include/b200_paged_attn.h).
//
// Argument validation mirrors the host-side checks of the reference launchers
// (unsupported head size / block size / dtype -> error before launch:
// paged_attention_v1.cu:122,156, quant_utils.cuh:554,580 of the reference); the
// kernels never report errors from the device and never synchronise the host.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace b200pa {

// kernels (other translation units)
int launch_reshape_and_cache(const void*, const void*, void*, void*, const int64_t*, int, int64_t,
                             int64_t, int, int, int, int, int, int, const float*, const float*,
                             cudaStream_t);
int launch_reshape_and_cache_flash(const void*, const void*, void*, void*, const int64_t*, int,
                                   int64_t, int64_t, int64_t, int64_t, int64_t, int, int, int, int,
                                   int, const float*, const float*, cudaStream_t);
int launch_convert_fp8(void*, const void*, int64_t, float, int, int, int, cudaStream_t);
int launch_reduce(void*, const float*, const float*, const void*, const int32_t*, int, int, int,
                  int, int, cudaStream_t);
int launch_generic(void*, float*, int64_t, int64_t, float*, float*, void*, const void*, const void*, const void*, int, int,
                   int, int, int, float, const int32_t*, int, const int32_t*, int, int,
                   const float*, int64_t, int64_t, int64_t, int, int, const float*, const float*,
                   BlockSparse, bool, cudaStream_t);
bool stream_supported(const void*, const void*, const void*, int, int, int, int, int64_t, int64_t,
                      int64_t, int, int, int);
int launch_stream(void*, float*, int64_t, int64_t, void*, int64_t, const void*, const void*, const void*, int, int, int, int,
                  float, const int32_t*, int, const int32_t*, const float*, int64_t, int64_t,
                  int64_t, int, int, const float*, const float*, const FusedWrite*,
                  const BlockSparse*, int64_t, int, int, int, cudaStream_t);
bool stream_flash_supported(const void*, const void*, const void*, int, int, int, int, int64_t,
                            int64_t, int64_t, int64_t, int, int);
bool stream_fused_supported(const FusedWrite&, int);
int64_t stream_workspace_bytes(int, int);
int launch_merge_attn_states(void*, float*, const void*, const float*, const void*, const float*,
                             int64_t, int, int, int, cudaStream_t);
int launch_correct_attn_out(void*, float*, const float*, int, int, int, int, int, int64_t, int64_t,
                            int64_t, int64_t, int64_t, int, cudaStream_t);

int launch_copy_blocks(void* const*, void* const*, int, const int64_t*, int, int64_t, cudaStream_t);
int launch_gather_cache(const void*, void*, const int32_t*, const int32_t*, int, int, int, int64_t,
                        int64_t, int64_t, int64_t, int, int, const float*, const int32_t*, int64_t,
                        int, cudaStream_t);
int launch_cp_gather_cache(const void*, void*, const int32_t*, const int32_t*, int, int, int,
                           int64_t, int64_t, int64_t, int64_t, int, const int32_t*, int64_t, int,
                           cudaStream_t);

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int check_launch(const char* what) {
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess)
    return set_error(B200PA_ERR_CUDA, "%s: CUDA error: %s", what, cudaGetErrorString(e));
  return B200PA_OK;
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

// SM count of the CURRENT device (the persistent grid is sized from it), cached per ordinal.
static int num_sms() {
  static std::atomic<int> cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      return 148;
    cached[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

// test hooks (include/b200_paged_attn_testing.h); constants in a build without them
#if B200PA_TESTING
static std::atomic<int> g_impl{B200PA_IMPL_AUTO};
static std::atomic<int> g_tune{-1};
static std::atomic<void*> g_trace{nullptr};
static int current_impl() { return g_impl.load(std::memory_order_relaxed); }
static int stream_tune() {
  int t = g_tune.load(std::memory_order_relaxed);
  if (t < 0) {
    const char* e = getenv("B200PA_STREAM_TUNE");
    t = e ? atoi(e) : 0;
    if (t < 0) t = 0;
    g_tune.store(t, std::memory_order_relaxed);
  }
  return t;
}
void* testing_trace_buffer() { return g_trace.load(std::memory_order_relaxed); }
#else
static int current_impl() { return B200PA_IMPL_AUTO; }
static int stream_tune() { return 0; }
void* testing_trace_buffer() { return nullptr; }
#endif

static bool head_size_ok(int h) {
  switch (h) {  // paged_attention_v1.cu:90-124 of the reference
    case 32: case 64: case 80: case 96: case 112: case 120: case 128: case 192: case 256: return true;
    default: return false;
  }
}

static int paged_attention_common(
    bool partitioned, void* out, float* lse, int64_t lse_seq_stride, int64_t lse_head_stride,
    float* exp_sums, float* max_logits, void* tmp_out,
    const void* query, const void* key_cache, const void* value_cache, int num_seqs,
    int num_heads, int num_kv_heads, int head_size, int block_size, float scale,
    const int32_t* block_tables, int max_num_blocks_per_seq, const int32_t* seq_lens,
    int max_seq_len, int max_num_partitions, const float* alibi_slopes, int64_t q_stride,
    int64_t kv_block_stride, int64_t kv_head_stride, int scalar_type, int kv_dtype,
    const float* k_scale, const float* v_scale, int tp_rank, int bs_local_blocks,
    int bs_vert_stride, int bs_block_size, int bs_head_sliding_step, void* workspace,
    int64_t workspace_bytes, void* stream_v, const FusedWrite* fused = nullptr) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  if (!head_size_ok(head_size))
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported head size: %d", head_size);
  if (block_size != 8 && block_size != 16 && block_size != 32)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported block size: %d", block_size);
  if (scalar_type != B200PA_F32 && scalar_type != B200PA_F16 && scalar_type != B200PA_BF16)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported input type of kv cache: %d", scalar_type);
  if (kv_dtype != B200PA_KV_AUTO && kv_dtype != B200PA_KV_FP8_E4M3 && kv_dtype != B200PA_KV_FP8_E5M2)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported data type of kv cache: %d", kv_dtype);
  if (num_seqs < 0 || num_heads <= 0 || num_kv_heads <= 0 || num_heads % num_kv_heads != 0)
    return set_error(B200PA_ERR_INVALID, "num_heads %d must be a positive multiple of num_kv_heads %d",
                     num_heads, num_kv_heads);
  if (!out || !query || !key_cache || !value_cache || !block_tables || !seq_lens)
    return set_error(B200PA_ERR_INVALID, "null tensor pointer");
  if (kv_dtype != B200PA_KV_AUTO && (!k_scale || !v_scale))
    return set_error(B200PA_ERR_INVALID, "fp8 kv cache needs k_scale and v_scale");
  const int ces = kv_dtype == B200PA_KV_AUTO ? (scalar_type == B200PA_F32 ? 4 : 2) : 1;
  const int x = 16 / ces;
  if (head_size % x != 0)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported head size: %d with %d-byte cache elements",
                     head_size, ces);
  if ((reinterpret_cast<uintptr_t>(key_cache) | reinterpret_cast<uintptr_t>(value_cache)) & 15u ||
      (kv_block_stride * ces) % 16 != 0 || (kv_head_stride * ces) % 16 != 0)
    return set_error(B200PA_ERR_INVALID, "kv cache must be 16-byte aligned (base and strides)");
  if (partitioned) {
    if (!exp_sums || !max_logits || !tmp_out)
      return set_error(B200PA_ERR_INVALID, "paged_attention_v2 needs exp_sums, max_logits, tmp_out");
    const int parts = (max_seq_len + kPartitionSize - 1) / kPartitionSize;
    if (max_num_partitions < parts)
      return set_error(B200PA_ERR_INVALID, "max_num_partitions %d < ceil(max_seq_len %d / 512)",
                       max_num_partitions, max_seq_len);
  }
  if (num_seqs == 0) return B200PA_OK;

  BlockSparse bsp;
  bsp.enabled = bs_vert_stride > 1;  // paged_attention_v1.cu:175 of the reference
  bsp.local_blocks = bs_local_blocks;
  bsp.vert_stride = bs_vert_stride;
  bsp.block_size = bs_block_size;
  bsp.head_sliding_step = bs_head_sliding_step;
  bsp.tp_rank = tp_rank;
  if (bsp.enabled && bs_block_size <= 0)
    return set_error(B200PA_ERR_INVALID, "blocksparse_block_size must be positive");

  const int impl = current_impl();
  const bool stream_ok =
      stream_supported(query, key_cache, value_cache, num_heads, num_kv_heads, head_size,
                       block_size, q_stride, kv_block_stride, kv_head_stride, scalar_type,
                       kv_dtype, bs_vert_stride);
  if (impl == B200PA_IMPL_STREAM && !stream_ok)
    return set_error(B200PA_ERR_UNSUPPORTED, "configuration not covered by the stream kernel");
  const bool use_stream =
      stream_ok && ((impl == B200PA_IMPL_AUTO && workspace) || impl == B200PA_IMPL_STREAM);
  if (fused && !(use_stream && stream_fused_supported(*fused, scalar_type))) {
    // decode_step outside the fused kernel's domain: the two launches it stands for
    int rc = launch_reshape_and_cache(fused->key, fused->value, const_cast<void*>(key_cache),
                                      const_cast<void*>(value_cache), fused->slot_mapping, num_seqs,
                                      fused->key_stride, fused->value_stride, num_kv_heads,
                                      head_size, block_size, x, scalar_type, kv_dtype, k_scale,
                                      v_scale, stream);
    if (rc) return rc;
    fused = nullptr;
  }
  if (use_stream) {
    // persistent stream-K kernel + boundary merge; serves v1 and v2 alike (the caller's v2
    // scratch is left untouched)
    return launch_stream(out, lse, lse_seq_stride, lse_head_stride, workspace, workspace_bytes, query, key_cache, value_cache, num_seqs,
                         num_heads, num_kv_heads, head_size, scale, block_tables,
                         max_num_blocks_per_seq, seq_lens, alibi_slopes, q_stride,
                         kv_block_stride, kv_head_stride, scalar_type, kv_dtype, k_scale, v_scale,
                         fused, &bsp, 0, block_size, num_sms(), stream_tune(), stream);
  }
  int rc = launch_generic(out, lse, lse_seq_stride, lse_head_stride, exp_sums, max_logits, tmp_out, query, key_cache, value_cache,
                          num_seqs, num_heads, num_kv_heads, head_size, block_size, scale,
                          block_tables, max_num_blocks_per_seq, seq_lens, max_seq_len,
                          max_num_partitions, alibi_slopes, q_stride, kv_block_stride,
                          kv_head_stride, scalar_type, kv_dtype, k_scale, v_scale, bsp, partitioned,
                          stream);
  if (rc) return rc;
  if (partitioned)
    rc = launch_reduce(out, exp_sums, max_logits, tmp_out, seq_lens, num_seqs, num_heads,
                       head_size, max_num_partitions, scalar_type, stream);
  return rc;
}

}  // namespace b200pa

using namespace b200pa;

extern "C" {

int b200pa_abi_version(void) { return B200PA_ABI_VERSION; }
const char* b200pa_last_error(void) { return g_err; }

int b200pa_kv_dtype_from_string(const char* s) {
  if (!s) return -1;
  if (!strcmp(s, "auto")) return B200PA_KV_AUTO;
  if (!strcmp(s, "fp8") || !strcmp(s, "fp8_e4m3")) return B200PA_KV_FP8_E4M3;
  if (!strcmp(s, "fp8_e5m2")) return B200PA_KV_FP8_E5M2;
  return -1;
}

#if B200PA_TESTING
void b200pa_set_impl(int impl) { g_impl.store(impl); }
int b200pa_get_impl(void) { return g_impl.load(); }
void b200pa_testing_set_stream_tune(int tune) { g_tune.store(tune < 0 ? 0 : tune); }
void b200pa_testing_set_trace(void* device_buffer, int64_t bytes) {
  g_trace.store(bytes >= 256 ? device_buffer : nullptr);
}
#endif
uint64_t b200pa_launch_count(void) { return g_launches.load(); }

int64_t b200pa_paged_attention_workspace_bytes(int head_size) {
  return stream_workspace_bytes(num_sms(), head_size > 0 ? head_size : 128);
}

int b200pa_paged_attention_v1(
    void* out, const void* query, const void* key_cache, const void* value_cache, int num_seqs,
    int num_heads, int num_kv_heads, int head_size, int block_size, float scale,
    const int32_t* block_tables, int max_num_blocks_per_seq, const int32_t* seq_lens,
    int max_seq_len, const float* alibi_slopes, int64_t q_stride, int64_t kv_block_stride,
    int64_t kv_head_stride, int scalar_type, int kv_dtype, const float* k_scale,
    const float* v_scale, int tp_rank, int blocksparse_local_blocks, int blocksparse_vert_stride,
    int blocksparse_block_size, int blocksparse_head_sliding_step, void* workspace,
    int64_t workspace_bytes, void* stream) {
  return paged_attention_common(false, out, nullptr, 0, 0, nullptr, nullptr, nullptr, query, key_cache,
                                value_cache, num_seqs, num_heads, num_kv_heads, head_size,
                                block_size, scale, block_tables, max_num_blocks_per_seq, seq_lens,
                                max_seq_len, 1, alibi_slopes, q_stride, kv_block_stride,
                                kv_head_stride, scalar_type, kv_dtype, k_scale, v_scale, tp_rank,
                                blocksparse_local_blocks, blocksparse_vert_stride,
                                blocksparse_block_size, blocksparse_head_sliding_step, workspace,
                                workspace_bytes, stream);
}

int b200pa_paged_attention_v2(
    void* out, float* exp_sums, float* max_logits, void* tmp_out, const void* query,
    const void* key_cache, const void* value_cache, int num_seqs, int num_heads, int num_kv_heads,
    int head_size, int block_size, float scale, const int32_t* block_tables,
    int max_num_blocks_per_seq, const int32_t* seq_lens, int max_seq_len, int max_num_partitions,
    const float* alibi_slopes, int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale, int tp_rank,
    int blocksparse_local_blocks, int blocksparse_vert_stride, int blocksparse_block_size,
    int blocksparse_head_sliding_step, void* workspace, int64_t workspace_bytes, void* stream) {
  return paged_attention_common(true, out, nullptr, 0, 0, exp_sums, max_logits, tmp_out, query, key_cache,
                                value_cache, num_seqs, num_heads, num_kv_heads, head_size,
                                block_size, scale, block_tables, max_num_blocks_per_seq, seq_lens,
                                max_seq_len, max_num_partitions, alibi_slopes, q_stride,
                                kv_block_stride, kv_head_stride, scalar_type, kv_dtype, k_scale,
                                v_scale, tp_rank, blocksparse_local_blocks,
                                blocksparse_vert_stride, blocksparse_block_size,
                                blocksparse_head_sliding_step, workspace, workspace_bytes, stream);
}

int b200pa_paged_attention_lse(
    void* out, float* lse, int64_t lse_seq_stride, int64_t lse_head_stride, const void* query,
    const void* key_cache, const void* value_cache, int num_seqs, int num_heads, int num_kv_heads,
    int head_size, int block_size, float scale, const int32_t* block_tables,
    int max_num_blocks_per_seq, const int32_t* seq_lens, int max_seq_len,
    const float* alibi_slopes, int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale, void* workspace,
    int64_t workspace_bytes, void* stream) {
  if (!lse) return set_error(B200PA_ERR_INVALID, "paged_attention_lse: null lse pointer");
  return paged_attention_common(false, out, lse, lse_seq_stride, lse_head_stride, nullptr, nullptr,
                                nullptr, query, key_cache, value_cache, num_seqs, num_heads,
                                num_kv_heads, head_size, block_size, scale, block_tables,
                                max_num_blocks_per_seq, seq_lens, max_seq_len, 1, alibi_slopes,
                                q_stride, kv_block_stride, kv_head_stride, scalar_type, kv_dtype,
                                k_scale, v_scale, 0, 0, 0, 64, 0, workspace, workspace_bytes,
                                stream);
}

int b200pa_decode_step(
    void* out, float* lse, int64_t lse_seq_stride, int64_t lse_head_stride, const void* query,
    const void* key, const void* value, int64_t key_stride, int64_t value_stride,
    const int64_t* slot_mapping, void* key_cache, void* value_cache, int num_seqs, int num_heads,
    int num_kv_heads, int head_size, int block_size, float scale, const int32_t* block_tables,
    int max_num_blocks_per_seq, const int32_t* seq_lens, int max_seq_len,
    const float* alibi_slopes, int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale, void* workspace,
    int64_t workspace_bytes, void* stream) {
  if (!key || !value || !slot_mapping)
    return set_error(B200PA_ERR_INVALID, "decode_step: null key / value / slot_mapping pointer");
  // reshape_and_cache indexes a dense cache (cache_kernels.cu:231-247 of the reference)
  if (kv_head_stride != (int64_t)head_size * block_size ||
      kv_block_stride != (int64_t)num_kv_heads * head_size * block_size)
    return set_error(B200PA_ERR_INVALID, "decode_step: the kv cache must be dense");
  const FusedWrite fw{key, value, key_stride, value_stride, slot_mapping};
  return paged_attention_common(false, out, lse, lse_seq_stride, lse_head_stride, nullptr, nullptr,
                                nullptr, query, key_cache, value_cache, num_seqs, num_heads,
                                num_kv_heads, head_size, block_size, scale, block_tables,
                                max_num_blocks_per_seq, seq_lens, max_seq_len, 1, alibi_slopes,
                                q_stride, kv_block_stride, kv_head_stride, scalar_type, kv_dtype,
                                k_scale, v_scale, 0, 0, 0, 64, 0, workspace, workspace_bytes,
                                stream, &fw);
}

int b200pa_paged_attention_flash(
    void* out, float* lse, int64_t lse_seq_stride, int64_t lse_head_stride, const void* query,
    const void* key, const void* value, int64_t key_stride, int64_t value_stride,
    const int64_t* slot_mapping, void* key_cache, void* value_cache, int num_seqs, int num_heads,
    int num_kv_heads, int head_size, int block_size, float scale, const int32_t* block_tables,
    int max_num_blocks_per_seq, const int32_t* seq_lens, int max_seq_len,
    const float* alibi_slopes, int64_t q_stride, int64_t block_stride, int64_t page_stride,
    int64_t head_stride, int scalar_type, int kv_dtype, const float* k_scale,
    const float* v_scale, void* workspace, int64_t workspace_bytes, void* stream_v) {
  (void)max_seq_len;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
  if (!out || !query || !key_cache || !value_cache || !block_tables || !seq_lens)
    return set_error(B200PA_ERR_INVALID, "null tensor pointer");
  if (num_seqs < 0 || num_heads <= 0 || num_kv_heads <= 0 || num_heads % num_kv_heads != 0)
    return set_error(B200PA_ERR_INVALID, "num_heads %d must be a positive multiple of num_kv_heads %d",
                     num_heads, num_kv_heads);
  if (kv_dtype != B200PA_KV_AUTO && (!k_scale || !v_scale))
    return set_error(B200PA_ERR_INVALID, "fp8 kv cache needs k_scale and v_scale");
  if (head_size != 32 && head_size != 64 && head_size != 80 && head_size != 96 &&
      head_size != 112 && head_size != 128 && head_size != 192 && head_size != 256)
    return set_error(B200PA_ERR_UNSUPPORTED,
                     "Unsupported head size: %d (flash layout: 32, 64, 80, 96, 112, 128, 192, 256)",
                     head_size);
  if (block_size != 16 && block_size != 32)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported block size: %d (flash layout: 16, 32)",
                     block_size);
  if (!stream_flash_supported(query, key_cache, value_cache, num_heads, num_kv_heads, head_size,
                              block_size, q_stride, block_stride, page_stride, head_stride,
                              scalar_type, kv_dtype))
    return set_error(B200PA_ERR_UNSUPPORTED,
                     "paged_attention_flash: needs a 16-bit query, an auto or fp8 cache and "
                     "16-byte aligned cache rows");
  const bool want_write = key || value || slot_mapping;
  if (want_write && (!key || !value || !slot_mapping))
    return set_error(B200PA_ERR_INVALID, "paged_attention_flash: key, value and slot_mapping go together");
  if (num_seqs == 0) return B200PA_OK;
  FusedWrite fw{key, value, key_stride, value_stride, slot_mapping};
  const FusedWrite* fused = want_write ? &fw : nullptr;
  if (fused && !stream_fused_supported(fw, scalar_type)) {
    // rows the fused path cannot load: the write it stands for, as its own launch
    int rc = launch_reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping,
                                            num_seqs, key_stride, value_stride, block_stride,
                                            page_stride, head_stride, num_kv_heads, head_size,
                                            block_size, scalar_type, kv_dtype, k_scale, v_scale,
                                            stream);
    if (rc) return rc;
    fused = nullptr;
  }
  return launch_stream(out, lse, lse_seq_stride, lse_head_stride, workspace, workspace_bytes, query,
                       key_cache, value_cache, num_seqs, num_heads, num_kv_heads, head_size, scale,
                       block_tables, max_num_blocks_per_seq, seq_lens, alibi_slopes, q_stride,
                       block_stride, head_stride, scalar_type, kv_dtype, k_scale, v_scale, fused,
                       nullptr, page_stride, block_size, num_sms(), stream_tune(), stream);
}

int b200pa_merge_attn_states(void* output, float* output_lse, const void* prefix_output,
                             const float* prefix_lse, const void* suffix_output,
                             const float* suffix_lse, int64_t num_tokens, int num_heads,
                             int head_size, int scalar_type, void* stream) {
  if (num_tokens < 0 || num_heads <= 0 || head_size <= 0)
    return set_error(B200PA_ERR_INVALID, "merge_attn_states: bad shape");
  if (scalar_type != B200PA_F32 && scalar_type != B200PA_F16 && scalar_type != B200PA_BF16)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported data type of O: %d", scalar_type);
  const int pack = scalar_type == B200PA_F32 ? 4 : 8;
  if (head_size % pack != 0)
    return set_error(B200PA_ERR_UNSUPPORTED, "headsize must be multiple of pack_size:%d", pack);
  if (num_tokens == 0) return B200PA_OK;
  if (!output || !prefix_output || !prefix_lse || !suffix_output || !suffix_lse)
    return set_error(B200PA_ERR_INVALID, "merge_attn_states: null tensor pointer");
  if ((reinterpret_cast<uintptr_t>(output) | reinterpret_cast<uintptr_t>(prefix_output) |
       reinterpret_cast<uintptr_t>(suffix_output)) & 15u)
    return set_error(B200PA_ERR_INVALID, "merge_attn_states: outputs must be 16-byte aligned");
  return launch_merge_attn_states(output, output_lse, prefix_output, prefix_lse, suffix_output,
                                  suffix_lse, num_tokens, num_heads, head_size, scalar_type,
                                  reinterpret_cast<cudaStream_t>(stream));
}

int b200pa_correct_attn_out(void* out, float* lse_out, const float* lses, int num_ranks,
                            int rank, int num_seqs, int num_heads, int head_size,
                            int64_t out_seq_stride, int64_t out_head_stride,
                            int64_t lses_rank_stride, int64_t lses_seq_stride,
                            int64_t lses_head_stride, int scalar_type, void* stream) {
  if (num_ranks <= 0 || rank < 0 || rank >= num_ranks || num_seqs < 0 || num_heads <= 0 ||
      head_size <= 0)
    return set_error(B200PA_ERR_INVALID, "correct_attn_out: bad shape");
  if (scalar_type != B200PA_F32 && scalar_type != B200PA_F16 && scalar_type != B200PA_BF16)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported data type of O: %d", scalar_type);
  if (num_seqs == 0) return B200PA_OK;
  if (!out || !lse_out || !lses)
    return set_error(B200PA_ERR_INVALID, "correct_attn_out: null tensor pointer");
  return launch_correct_attn_out(out, lse_out, lses, num_ranks, rank, num_seqs, num_heads,
                                 head_size, out_seq_stride, out_head_stride, lses_rank_stride,
                                 lses_seq_stride, lses_head_stride, scalar_type,
                                 reinterpret_cast<cudaStream_t>(stream));
}

static int check_cache_args(const char* op, const void* key, const void* value, void* kc, void* vc,
                            const int64_t* slots, int num_tokens, int scalar_type, int kv_dtype,
                            const float* ks, const float* vs) {
  if (num_tokens < 0) return set_error(B200PA_ERR_INVALID, "%s: negative num_tokens", op);
  if (num_tokens > 0 && (!key || !value || !kc || !vc || !slots))
    return set_error(B200PA_ERR_INVALID, "%s: null tensor pointer", op);
  if (scalar_type != B200PA_F32 && scalar_type != B200PA_F16 && scalar_type != B200PA_BF16)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported input type of kv cache: %d", scalar_type);
  if (kv_dtype != B200PA_KV_AUTO && kv_dtype != B200PA_KV_FP8_E4M3 && kv_dtype != B200PA_KV_FP8_E5M2)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported data type of kv cache: %d", kv_dtype);
  if (kv_dtype != B200PA_KV_AUTO && (!ks || !vs))
    return set_error(B200PA_ERR_INVALID, "%s: fp8 kv cache needs k_scale and v_scale", op);
  return B200PA_OK;
}

int b200pa_reshape_and_cache(const void* key, const void* value, void* key_cache,
                             void* value_cache, const int64_t* slot_mapping, int num_tokens,
                             int64_t key_stride, int64_t value_stride, int num_heads,
                             int head_size, int block_size, int x, int scalar_type, int kv_dtype,
                             const float* k_scale, const float* v_scale, void* stream) {
  int rc = check_cache_args("reshape_and_cache", key, value, key_cache, value_cache, slot_mapping,
                            num_tokens, scalar_type, kv_dtype, k_scale, v_scale);
  if (rc) return rc;
  if (num_heads <= 0 || head_size <= 0 || block_size <= 0)
    return set_error(B200PA_ERR_INVALID, "reshape_and_cache: bad shape");
  return launch_reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, num_tokens,
                                  key_stride, value_stride, num_heads, head_size, block_size, x,
                                  scalar_type, kv_dtype, k_scale, v_scale,
                                  reinterpret_cast<cudaStream_t>(stream));
}

int b200pa_reshape_and_cache_flash(const void* key, const void* value, void* key_cache,
                                   void* value_cache, const int64_t* slot_mapping, int num_tokens,
                                   int64_t key_stride, int64_t value_stride, int64_t block_stride,
                                   int64_t page_stride, int64_t head_stride, int num_heads,
                                   int head_size, int block_size, int scalar_type, int kv_dtype,
                                   const float* k_scale, const float* v_scale, void* stream) {
  int rc = check_cache_args("reshape_and_cache_flash", key, value, key_cache, value_cache,
                            slot_mapping, num_tokens, scalar_type, kv_dtype, k_scale, v_scale);
  if (rc) return rc;
  if (num_heads <= 0 || head_size <= 0 || block_size <= 0)
    return set_error(B200PA_ERR_INVALID, "reshape_and_cache_flash: bad shape");
  return launch_reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping,
                                        num_tokens, key_stride, value_stride, block_stride,
                                        page_stride, head_stride, num_heads, head_size, block_size,
                                        scalar_type, kv_dtype, k_scale, v_scale,
                                        reinterpret_cast<cudaStream_t>(stream));
}

int b200pa_convert_fp8(void* dst, const void* src, int64_t num_elements, float scale,
                       int scalar_type, int kv_dtype, int to_fp8, void* stream) {
  if (num_elements < 0) return set_error(B200PA_ERR_INVALID, "convert_fp8: negative size");
  if (num_elements > 0 && (!dst || !src)) return set_error(B200PA_ERR_INVALID, "convert_fp8: null pointer");
  if (scalar_type != B200PA_F32 && scalar_type != B200PA_F16 && scalar_type != B200PA_BF16)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported input type of kv cache: %d", scalar_type);
  return launch_convert_fp8(dst, src, num_elements, scale, scalar_type, kv_dtype, to_fp8,
                            reinterpret_cast<cudaStream_t>(stream));
}

int b200pa_swap_blocks(const void* src, void* dst, int64_t block_bytes,
                       const int64_t* block_mapping_host, int64_t num_pairs, int direction,
                       void* stream) {
  if (num_pairs < 0 || block_bytes < 0) return set_error(B200PA_ERR_INVALID, "swap_blocks: negative size");
  if (num_pairs == 0 || block_bytes == 0) return B200PA_OK;
  if (!src || !dst || !block_mapping_host)
    return set_error(B200PA_ERR_INVALID, "swap_blocks: null pointer");
  cudaMemcpyKind kind;
  switch (direction) {
    case B200PA_COPY_D2D: kind = cudaMemcpyDeviceToDevice; break;
    case B200PA_COPY_D2H: kind = cudaMemcpyDeviceToHost; break;
    case B200PA_COPY_H2D: kind = cudaMemcpyHostToDevice; break;
    default: return set_error(B200PA_ERR_INVALID, "Invalid device combination");
  }
  const char* s = static_cast<const char*>(src);
  char* d = static_cast<char*>(dst);
  int64_t i = 0;
  while (i < num_pairs) {
    // one copy per run of consecutive (src + 1, dst + 1) pairs
    const int64_t s0 = block_mapping_host[2 * i], d0 = block_mapping_host[2 * i + 1];
    int64_t n = 1;
    while (i + n < num_pairs && block_mapping_host[2 * (i + n)] == s0 + n &&
           block_mapping_host[2 * (i + n) + 1] == d0 + n)
      ++n;
    const cudaError_t e = cudaMemcpyAsync(d + d0 * block_bytes, s + s0 * block_bytes,
                                          (size_t)(n * block_bytes), kind,
                                          reinterpret_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      return set_error(B200PA_ERR_CUDA, "swap_blocks: cudaMemcpyAsync: %s", cudaGetErrorString(e));
    }
    i += n;
  }
  return B200PA_OK;
}

int b200pa_copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs, int num_layers,
                       const int64_t* block_mapping, int num_pairs, int64_t block_bytes,
                       void* stream) {
  if (num_layers < 0 || num_pairs < 0 || block_bytes < 0)
    return set_error(B200PA_ERR_INVALID, "copy_blocks: negative size");
  if (num_layers == 0 || num_pairs == 0 || block_bytes == 0) return B200PA_OK;
  if (!key_cache_ptrs || !value_cache_ptrs || !block_mapping)
    return set_error(B200PA_ERR_INVALID, "copy_blocks: null pointer");
  for (int l = 0; l < num_layers; ++l)
    if (!key_cache_ptrs[l] || !value_cache_ptrs[l])
      return set_error(B200PA_ERR_INVALID, "copy_blocks: null cache pointer in layer %d", l);
  return launch_copy_blocks(key_cache_ptrs, value_cache_ptrs, num_layers, block_mapping, num_pairs,
                            block_bytes, reinterpret_cast<cudaStream_t>(stream));
}

static int check_gather_args(const char* op, const void* src_cache, void* dst,
                             const int32_t* block_table, const int32_t* cu_seq_lens,
                             int batch_size, int block_size, int entry_size) {
  if (batch_size < 0) return set_error(B200PA_ERR_INVALID, "%s: negative batch_size", op);
  if (block_size <= 0 || entry_size <= 0) return set_error(B200PA_ERR_INVALID, "%s: bad shape", op);
  if (!cu_seq_lens) return set_error(B200PA_ERR_INVALID, "%s: null cu_seq_lens", op);
  if (batch_size > 0 && (!src_cache || !dst || !block_table))
    return set_error(B200PA_ERR_INVALID, "%s: null tensor pointer", op);
  return B200PA_OK;
}

int b200pa_gather_and_maybe_dequant_cache(
    const void* src_cache, void* dst, const int32_t* block_table, const int32_t* cu_seq_lens,
    int batch_size, int block_size, int entry_size, int64_t block_table_stride,
    int64_t cache_block_stride, int64_t cache_entry_stride, int64_t dst_entry_stride,
    int scalar_type, int kv_dtype, const float* scale, const int32_t* seq_starts,
    int64_t tot_tokens, void* stream) {
  int rc = check_gather_args("gather_and_maybe_dequant_cache", src_cache, dst, block_table,
                             cu_seq_lens, batch_size, block_size, entry_size);
  if (rc) return rc;
  if (scalar_type != B200PA_F32 && scalar_type != B200PA_F16 && scalar_type != B200PA_BF16)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported input type of kv cache: %d", scalar_type);
  if (kv_dtype != B200PA_KV_AUTO && kv_dtype != B200PA_KV_FP8_E4M3 && kv_dtype != B200PA_KV_FP8_E5M2)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported data type of kv cache: %d", kv_dtype);
  if (kv_dtype != B200PA_KV_AUTO && !scale)
    return set_error(B200PA_ERR_INVALID, "gather_and_maybe_dequant_cache: fp8 cache needs scale");
  if (batch_size == 0 || tot_tokens == 0) return B200PA_OK;
  return launch_gather_cache(src_cache, dst, block_table, cu_seq_lens, batch_size, block_size,
                             entry_size, block_table_stride, cache_block_stride, cache_entry_stride,
                             dst_entry_stride, scalar_type, kv_dtype, scale, seq_starts, tot_tokens,
                             num_sms(), reinterpret_cast<cudaStream_t>(stream));
}

int b200pa_cp_gather_cache(
    const void* src_cache, void* dst, const int32_t* block_table, const int32_t* cu_seq_lens,
    int batch_size, int block_size, int entry_size, int64_t block_table_stride,
    int64_t cache_block_stride, int64_t cache_entry_stride, int64_t dst_entry_stride,
    int element_bytes, const int32_t* seq_starts, int64_t tot_tokens, void* stream) {
  int rc = check_gather_args("cp_gather_cache", src_cache, dst, block_table, cu_seq_lens,
                             batch_size, block_size, entry_size);
  if (rc) return rc;
  if (element_bytes != 1 && element_bytes != 2 && element_bytes != 4)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported data type width: %d", element_bytes * 8);
  if (batch_size == 0 || tot_tokens == 0) return B200PA_OK;
  return launch_cp_gather_cache(src_cache, dst, block_table, cu_seq_lens, batch_size, block_size,
                                entry_size, block_table_stride, cache_block_stride,
                                cache_entry_stride, dst_entry_stride, element_bytes, seq_starts,
                                tot_tokens, num_sms(), reinterpret_cast<cudaStream_t>(stream));
}

int b200pa_get_max_shared_memory_per_block_device_attribute(int device_id, int64_t* value) {
  if (!value) return set_error(B200PA_ERR_INVALID, "null output pointer");
  int v = 0;
  const cudaError_t e = cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, device_id);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return set_error(B200PA_ERR_CUDA, "cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
  }
  *value = v;
  return B200PA_OK;
}

}  // extern "C"
