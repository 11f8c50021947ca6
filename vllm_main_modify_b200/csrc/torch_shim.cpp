// Thin TORCH_LIBRARY shim over the C ABI (include/b200_paged_attn.h).
//
// Exports, VERBATIM, the op schemas of the reference
// (csrc/torch_bindings.cpp:52-76 `_C`, :667-699 `_C_cache_ops`, :718-730
// `_C_cuda_utils`) with the C++ signatures of csrc/ops.h:30-52 and
// csrc/cache.h:20-41.  Each op only extracts pointers / sizes / strides from the
// tensors -- exactly the values the reference launchers read
// (paged_attention_v1.cu:54-75, cache_kernels.cu:426-433, :478-488) -- calls the
// C ABI on at::cuda::getCurrentCUDAStream() under an OptionalCUDAGuard, and turns
// a non-zero status into TORCH_CHECK(false, msg).
//
// The ops are always registered under the private namespace `b200pa`;
// b200pa_register_vllm_namespaces() additionally defines them as `_C::*`,
// `_C_cache_ops::*`, `_C_cuda_utils::*` (when the reference's own extension is
// not loaded in the process) so that an unmodified `vllm/_custom_ops.py` resolves
// `torch.ops._C.paged_attention_v1` etc. to this library.
#include <cstdlib>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGraphsC10Utils.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/all.h>
#include <torch/library.h>

#include <memory>
#include <optional>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "b200_paged_attn.h"
#ifndef B200PA_TESTING
#define B200PA_TESTING 0
#endif
#if B200PA_TESTING
#include "b200_paged_attn_testing.h"
#endif

namespace {

int scalar_code(at::ScalarType t) {
  switch (t) {
    case at::ScalarType::Float: return B200PA_F32;
    case at::ScalarType::Half: return B200PA_F16;
    case at::ScalarType::BFloat16: return B200PA_BF16;
    default: TORCH_CHECK(false, "Unsupported input type of kv cache: ", t);
  }
}

int kv_code(const std::string& s) {
  const int k = b200pa_kv_dtype_from_string(s.c_str());
  TORCH_CHECK(k >= 0, "Unsupported data type of kv cache: ", s);
  return k;
}

void check_status(int rc) { TORCH_CHECK(rc == 0, b200pa_last_error()); }

void check_cuda(const torch::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor (this library has no CPU path)");
}

void check_cache_dtype(const torch::Tensor& cache, const torch::Tensor& ref, int kv,
                       const char* name) {
  if (kv == B200PA_KV_AUTO) {
    TORCH_CHECK(cache.scalar_type() == ref.scalar_type(), name,
                " dtype must match the query/key dtype when kv_cache_dtype is auto");
  } else {
    TORCH_CHECK(cache.element_size() == 1, name, " must be a 1-byte tensor for an fp8 kv cache");
  }
}

// Scratch of the stream kernel: ONE workspace per (device, stream), allocated on first use,
// zero-filled once (the ticket counters in its header, include/b200_paged_attn.h) and kept -- no
// allocator round trip per call, and a CUDA-graph capture sees the same pointer in every replay.
// A call on another stream (a capture stream included) gets its own workspace, so concurrent
// streams never share counters or slots.  Sized for the largest head size (256), so every shape
// can use it.
// The allocation and the zero fill always EXECUTE before this function returns the tensor's first
// use: when the first use of a stream falls inside a CUDA-graph capture, they are made on a side
// stream that is not capturing (ordinary allocator pool, not the graph's private one) and waited
// for -- a memset merely recorded into that graph would leave the counters uninitialised for every
// eager call and every other graph on the stream until that one graph is replayed.
// The map is leaked on purpose: no tensor destructor may run after CUDA has been torn down at exit.
torch::Tensor cached_workspace(const torch::Tensor& query) {
  static std::mutex mu;
  static auto* cache = new std::map<std::pair<int, cudaStream_t>, torch::Tensor>();
  const int dev = query.get_device();
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream(dev);
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache->find({dev, stream});
  if (it != cache->end()) return it->second;
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const int64_t bytes = b200pa_paged_attention_workspace_bytes(256);
  const auto opts = query.options().dtype(at::kByte);
  torch::Tensor ws;
  if (c10::cuda::currentStreamCaptureStatusMayInitCtx() != c10::cuda::CaptureStatus::None) {
    const c10::cuda::CUDAStreamCaptureModeGuard relaxed(cudaStreamCaptureModeRelaxed);
    const at::cuda::CUDAStream side = at::cuda::getStreamFromPool(false, dev);
    {
      const at::cuda::CUDAStreamGuard on_side(side);
      ws = at::zeros({bytes}, opts);
    }
    side.synchronize();
  } else {
    ws = at::zeros({bytes}, opts);
  }
  cache->emplace(std::make_pair(dev, stream), ws);
  return ws;
}

torch::Tensor attention_workspace(const torch::Tensor& query) {
#if B200PA_TESTING
  // forced generic kernel (tests): it takes no workspace
  if (b200pa_get_impl() == B200PA_IMPL_GENERIC) return torch::Tensor();
#endif
  return cached_workspace(query);
}

void paged_attention_v1(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache,
                        torch::Tensor& value_cache, int64_t num_kv_heads, double scale,
                        torch::Tensor& block_tables, torch::Tensor& seq_lens, int64_t block_size,
                        int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                        torch::Tensor& v_scale, const int64_t tp_rank,
                        const int64_t blocksparse_local_blocks,
                        const int64_t blocksparse_vert_stride,
                        const int64_t blocksparse_block_size,
                        const int64_t blocksparse_head_sliding_step) {
  check_cuda(query, "query");
  const int kv = kv_code(kv_cache_dtype);
  const int st = scalar_code(query.scalar_type());
  check_cache_dtype(key_cache, query, kv, "key_cache");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int, "block_tables must be int32");
  TORCH_CHECK(seq_lens.scalar_type() == at::ScalarType::Int, "seq_lens must be int32");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == query.size(2),
              "query heads must be densely packed");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  const float* alibi = alibi_slopes ? reinterpret_cast<const float*>(alibi_slopes->data_ptr()) : nullptr;
  const torch::Tensor ws = attention_workspace(query);
  check_status(b200pa_paged_attention_v1(
      out.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
      (int)query.size(0), (int)query.size(1), (int)num_kv_heads, (int)query.size(2),
      (int)block_size, (float)scale, block_tables.data_ptr<int>(), (int)block_tables.size(1),
      seq_lens.data_ptr<int>(), (int)max_seq_len, alibi, query.stride(0), key_cache.stride(0),
      key_cache.stride(1), st, kv, reinterpret_cast<const float*>(k_scale.data_ptr()),
      reinterpret_cast<const float*>(v_scale.data_ptr()), (int)tp_rank,
      (int)blocksparse_local_blocks, (int)blocksparse_vert_stride, (int)blocksparse_block_size,
      (int)blocksparse_head_sliding_step, ws.defined() ? ws.data_ptr() : nullptr,
      ws.defined() ? ws.numel() : 0, stream));
}

void paged_attention_v2_impl(bool want_ref_scratch,
                        torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits,
                        torch::Tensor& tmp_out, torch::Tensor& query, torch::Tensor& key_cache,
                        torch::Tensor& value_cache, int64_t num_kv_heads, double scale,
                        torch::Tensor& block_tables, torch::Tensor& seq_lens, int64_t block_size,
                        int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                        torch::Tensor& v_scale, const int64_t tp_rank,
                        const int64_t blocksparse_local_blocks,
                        const int64_t blocksparse_vert_stride,
                        const int64_t blocksparse_block_size,
                        const int64_t blocksparse_head_sliding_step) {
  check_cuda(query, "query");
  const int kv = kv_code(kv_cache_dtype);
  const int st = scalar_code(query.scalar_type());
  check_cache_dtype(key_cache, query, kv, "key_cache");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int, "block_tables must be int32");
  TORCH_CHECK(seq_lens.scalar_type() == at::ScalarType::Int, "seq_lens must be int32");
  TORCH_CHECK(exp_sums.scalar_type() == at::ScalarType::Float &&
                  max_logits.scalar_type() == at::ScalarType::Float,
              "exp_sums / max_logits must be float32");
  TORCH_CHECK(out.is_contiguous() && exp_sums.is_contiguous() && max_logits.is_contiguous() &&
                  tmp_out.is_contiguous(),
              "out / exp_sums / max_logits / tmp_out must be contiguous");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == query.size(2),
              "query heads must be densely packed");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  const float* alibi = alibi_slopes ? reinterpret_cast<const float*>(alibi_slopes->data_ptr()) : nullptr;
  // The stream kernel splits the context by its own schedule and keeps its partials in the
  // workspace: exp_sums / max_logits / tmp_out stay untouched (the reference's callers treat them
  // as scratch, paged_attn.py:163-173).  want_ref_scratch serves a caller that reads them: the call
  // then takes the generic kernel, which leaves the reference's per-partition partials
  // (attention_kernels.cuh:345-351, 487) in them.
  const torch::Tensor ws = want_ref_scratch ? torch::Tensor() : attention_workspace(query);
  check_status(b200pa_paged_attention_v2(
      out.data_ptr(), exp_sums.data_ptr<float>(), max_logits.data_ptr<float>(), tmp_out.data_ptr(),
      query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), (int)query.size(0),
      (int)query.size(1), (int)num_kv_heads, (int)query.size(2), (int)block_size, (float)scale,
      block_tables.data_ptr<int>(), (int)block_tables.size(1), seq_lens.data_ptr<int>(),
      (int)max_seq_len, (int)exp_sums.size(-1), alibi, query.stride(0), key_cache.stride(0),
      key_cache.stride(1), st, kv, reinterpret_cast<const float*>(k_scale.data_ptr()),
      reinterpret_cast<const float*>(v_scale.data_ptr()), (int)tp_rank,
      (int)blocksparse_local_blocks, (int)blocksparse_vert_stride, (int)blocksparse_block_size,
      (int)blocksparse_head_sliding_step, ws.defined() ? ws.data_ptr() : nullptr,
      ws.defined() ? ws.numel() : 0, stream));
}

#define B200PA_V2_ARGS                                                                             \
  torch::Tensor &out, torch::Tensor &exp_sums, torch::Tensor &max_logits, torch::Tensor &tmp_out,  \
      torch::Tensor &query, torch::Tensor &key_cache, torch::Tensor &value_cache,                  \
      int64_t num_kv_heads, double scale, torch::Tensor &block_tables, torch::Tensor &seq_lens,    \
      int64_t block_size, int64_t max_seq_len, const std::optional<torch::Tensor>&alibi_slopes,    \
      const std::string &kv_cache_dtype, torch::Tensor &k_scale, torch::Tensor &v_scale,           \
      const int64_t tp_rank, const int64_t blocksparse_local_blocks,                               \
      const int64_t blocksparse_vert_stride, const int64_t blocksparse_block_size,                 \
      const int64_t blocksparse_head_sliding_step
#define B200PA_V2_PASS                                                                             \
  out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache, num_kv_heads, scale,          \
      block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale,      \
      v_scale, tp_rank, blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size, \
      blocksparse_head_sliding_step

// The reference's op.  Process-wide switch for an UNMODIFIED caller that reads the partials:
// B200PA_V2_REFERENCE_SCRATCH=1; a caller that can name the op it wants uses the per-call form
// below instead.
void paged_attention_v2(B200PA_V2_ARGS) {
  const char* ref_scratch = std::getenv("B200PA_V2_REFERENCE_SCRATCH");
  paged_attention_v2_impl(ref_scratch != nullptr && ref_scratch[0] == '1', B200PA_V2_PASS);
}

// b200pa::paged_attention_v2_partials -- same schema, and exp_sums / max_logits / tmp_out are
// filled with the reference's per-partition partials on THIS call, whatever the environment says.
void paged_attention_v2_partials(B200PA_V2_ARGS) { paged_attention_v2_impl(true, B200PA_V2_PASS); }
#undef B200PA_V2_ARGS
#undef B200PA_V2_PASS

// Decode + log-sum-exp (b200pa namespace only; the reference has no such op -- it is what
// AttentionImpl.can_return_lse_for_decode, abstract.py:115-138, asks of a backend).
// lse: [num_seqs, num_heads] fp32, any strides.
void paged_attention_lse(torch::Tensor& out, torch::Tensor& lse, torch::Tensor& query,
                         torch::Tensor& key_cache, torch::Tensor& value_cache,
                         int64_t num_kv_heads, double scale, torch::Tensor& block_tables,
                         torch::Tensor& seq_lens, int64_t block_size, int64_t max_seq_len,
                         const std::optional<torch::Tensor>& alibi_slopes,
                         const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                         torch::Tensor& v_scale) {
  check_cuda(query, "query");
  const int kv = kv_code(kv_cache_dtype);
  const int st = scalar_code(query.scalar_type());
  check_cache_dtype(key_cache, query, kv, "key_cache");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int, "block_tables must be int32");
  TORCH_CHECK(seq_lens.scalar_type() == at::ScalarType::Int, "seq_lens must be int32");
  TORCH_CHECK(lse.scalar_type() == at::ScalarType::Float && lse.dim() == 2 &&
                  lse.size(0) == query.size(0) && lse.size(1) == query.size(1),
              "lse must be float32 [num_seqs, num_heads]");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == query.size(2),
              "query heads must be densely packed");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  const float* alibi = alibi_slopes ? reinterpret_cast<const float*>(alibi_slopes->data_ptr()) : nullptr;
  const torch::Tensor ws = attention_workspace(query);
  check_status(b200pa_paged_attention_lse(
      out.data_ptr(), lse.data_ptr<float>(), lse.stride(0), lse.stride(1), query.data_ptr(),
      key_cache.data_ptr(), value_cache.data_ptr(), (int)query.size(0), (int)query.size(1),
      (int)num_kv_heads, (int)query.size(2), (int)block_size, (float)scale,
      block_tables.data_ptr<int>(), (int)block_tables.size(1), seq_lens.data_ptr<int>(),
      (int)max_seq_len, alibi, query.stride(0), key_cache.stride(0), key_cache.stride(1), st, kv,
      reinterpret_cast<const float*>(k_scale.data_ptr()),
      reinterpret_cast<const float*>(v_scale.data_ptr()), ws.defined() ? ws.data_ptr() : nullptr,
      ws.defined() ? ws.numel() : 0, stream));
}

// reshape_and_cache + paged_attention_v1 (+ lse) of a pure decode batch in one launch
// (b200pa_decode_step); the checks are those of the two ops it stands for.
void decode_step(torch::Tensor& out, const std::optional<torch::Tensor>& lse, torch::Tensor& query,
                 torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache,
                 torch::Tensor& value_cache, torch::Tensor& slot_mapping, int64_t num_kv_heads,
                 double scale, torch::Tensor& block_tables, torch::Tensor& seq_lens,
                 int64_t block_size, int64_t max_seq_len,
                 const std::optional<torch::Tensor>& alibi_slopes,
                 const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                 torch::Tensor& v_scale) {
  check_cuda(query, "query");
  check_cuda(key, "key");
  const int kv = kv_code(kv_cache_dtype);
  const int st = scalar_code(query.scalar_type());
  check_cache_dtype(key_cache, query, kv, "key_cache");
  check_cache_dtype(value_cache, query, kv, "value_cache");
  TORCH_CHECK(key.scalar_type() == query.scalar_type() && value.scalar_type() == query.scalar_type(),
              "key / value dtype must match query");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int, "block_tables must be int32");
  TORCH_CHECK(seq_lens.scalar_type() == at::ScalarType::Int, "seq_lens must be int32");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long, "slot_mapping must be int64");
  TORCH_CHECK(key.size(0) == query.size(0) && slot_mapping.size(0) == query.size(0) &&
                  value.size(0) == query.size(0),
              "decode_step takes one new token per sequence");
  TORCH_CHECK(key.size(1) == num_kv_heads && key.size(2) == query.size(2) &&
                  key.stride(2) == 1 && key.stride(1) == key.size(2) && value.stride(2) == 1 &&
                  value.stride(1) == value.size(2),
              "key / value must be [num_seqs, num_kv_heads, head_size] with dense heads");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == query.size(2),
              "query heads must be densely packed");
  float* lse_ptr = nullptr;
  int64_t ls0 = 0, ls1 = 0;
  if (lse.has_value()) {
    TORCH_CHECK(lse->scalar_type() == at::ScalarType::Float && lse->dim() == 2 &&
                    lse->size(0) == query.size(0) && lse->size(1) == query.size(1),
                "lse must be float32 [num_seqs, num_heads]");
    lse_ptr = lse->data_ptr<float>();
    ls0 = lse->stride(0);
    ls1 = lse->stride(1);
  }
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  const float* alibi = alibi_slopes ? reinterpret_cast<const float*>(alibi_slopes->data_ptr()) : nullptr;
  const torch::Tensor ws = attention_workspace(query);
  check_status(b200pa_decode_step(
      out.data_ptr(), lse_ptr, ls0, ls1, query.data_ptr(), key.data_ptr(), value.data_ptr(),
      key.stride(0), value.stride(0), slot_mapping.data_ptr<int64_t>(), key_cache.data_ptr(),
      value_cache.data_ptr(), (int)query.size(0), (int)query.size(1), (int)num_kv_heads,
      (int)query.size(2), (int)block_size, (float)scale, block_tables.data_ptr<int>(),
      (int)block_tables.size(1), seq_lens.data_ptr<int>(), (int)max_seq_len, alibi,
      query.stride(0), key_cache.stride(0), key_cache.stride(1), st, kv,
      reinterpret_cast<const float*>(k_scale.data_ptr()),
      reinterpret_cast<const float*>(v_scale.data_ptr()), ws.defined() ? ws.data_ptr() : nullptr,
      ws.defined() ? ws.numel() : 0, stream));
}

// Decode attention over the token-major cache of reshape_and_cache_flash:
// key_cache / value_cache [num_blocks, block_size, num_kv_heads, head_size], NHD or HND by strides
// (the logical shape the reference's flash backends use).  key / value / slot_mapping given:
// the cache write happens in the same launch (b200pa_paged_attention_flash).
void paged_attention_flash(torch::Tensor& out, const std::optional<torch::Tensor>& lse,
                           torch::Tensor& query, const std::optional<torch::Tensor>& key,
                           const std::optional<torch::Tensor>& value, torch::Tensor& key_cache,
                           torch::Tensor& value_cache,
                           const std::optional<torch::Tensor>& slot_mapping, double scale,
                           torch::Tensor& block_tables, torch::Tensor& seq_lens,
                           int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
                           const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                           torch::Tensor& v_scale) {
  check_cuda(query, "query");
  const int kv = kv_code(kv_cache_dtype);
  const int st = scalar_code(query.scalar_type());
  check_cache_dtype(key_cache, query, kv, "key_cache");
  check_cache_dtype(value_cache, query, kv, "value_cache");
  TORCH_CHECK(key_cache.dim() == 4 && value_cache.dim() == 4,
              "flash caches are [num_blocks, block_size, num_kv_heads, head_size]");
  TORCH_CHECK(key_cache.stride(3) == 1 && value_cache.stride(3) == 1,
              "head elements must be contiguous in the cache");
  TORCH_CHECK(key_cache.strides() == value_cache.strides(),
              "key_cache and value_cache must share their strides");
  TORCH_CHECK(key_cache.size(3) == query.size(2), "head_size of cache and query differ");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int, "block_tables must be int32");
  TORCH_CHECK(seq_lens.scalar_type() == at::ScalarType::Int, "seq_lens must be int32");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == query.size(2),
              "query heads must be densely packed");
  const int64_t num_kv_heads = key_cache.size(2);
  float* lse_ptr = nullptr;
  int64_t ls0 = 0, ls1 = 0;
  if (lse.has_value()) {
    TORCH_CHECK(lse->scalar_type() == at::ScalarType::Float && lse->dim() == 2 &&
                    lse->size(0) == query.size(0) && lse->size(1) == query.size(1),
                "lse must be float32 [num_seqs, num_heads]");
    lse_ptr = lse->data_ptr<float>();
    ls0 = lse->stride(0);
    ls1 = lse->stride(1);
  }
  const void *kp = nullptr, *vp = nullptr;
  const int64_t* sp = nullptr;
  int64_t kst = 0, vst = 0;
  TORCH_CHECK(key.has_value() == value.has_value() && key.has_value() == slot_mapping.has_value(),
              "key, value and slot_mapping go together");
  if (key.has_value()) {
    TORCH_CHECK(key->scalar_type() == query.scalar_type() &&
                    value->scalar_type() == query.scalar_type(),
                "key / value dtype must match query");
    TORCH_CHECK(slot_mapping->scalar_type() == at::ScalarType::Long, "slot_mapping must be int64");
    TORCH_CHECK(key->size(0) == query.size(0) && value->size(0) == query.size(0) &&
                    slot_mapping->size(0) == query.size(0),
                "the fused write takes one new token per sequence");
    TORCH_CHECK(key->size(1) == num_kv_heads && key->size(2) == query.size(2) &&
                    key->stride(2) == 1 && key->stride(1) == key->size(2) &&
                    value->stride(2) == 1 && value->stride(1) == value->size(2),
                "key / value must be [num_seqs, num_kv_heads, head_size] with dense heads");
    kp = key->data_ptr();
    vp = value->data_ptr();
    sp = slot_mapping->data_ptr<int64_t>();
    kst = key->stride(0);
    vst = value->stride(0);
  }
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  const float* alibi = alibi_slopes ? reinterpret_cast<const float*>(alibi_slopes->data_ptr()) : nullptr;
  const torch::Tensor ws = cached_workspace(query);
  check_status(b200pa_paged_attention_flash(
      out.data_ptr(), lse_ptr, ls0, ls1, query.data_ptr(), kp, vp, kst, vst, sp,
      key_cache.data_ptr(), value_cache.data_ptr(), (int)query.size(0), (int)query.size(1),
      (int)num_kv_heads, (int)query.size(2), (int)key_cache.size(1), (float)scale,
      block_tables.data_ptr<int>(), (int)block_tables.size(1), seq_lens.data_ptr<int>(),
      (int)max_seq_len, alibi, query.stride(0), key_cache.stride(0), key_cache.stride(1),
      key_cache.stride(2), st, kv, reinterpret_cast<const float*>(k_scale.data_ptr()),
      reinterpret_cast<const float*>(v_scale.data_ptr()), ws.data_ptr(), ws.numel(), stream));
}

// csrc/attention/merge_attn_states.cu:131-180 of the reference (same checks, same messages)
// Causal attention of query chunks over the paged cache (b200pa_prefill_attention): what the
// reference's paged backends call flash_attn_varlen_func(..., block_table) / context_attention_fwd
// for (cpu_attn.py:584-612, paged_attn.py:200-242).  heads_per_tile = 0: the library chooses.
void prefill_attention(torch::Tensor& out, const std::optional<torch::Tensor>& lse,
                       torch::Tensor& query, torch::Tensor& key_cache, torch::Tensor& value_cache,
                       double scale, torch::Tensor& block_tables, torch::Tensor& query_start_loc,
                       torch::Tensor& seq_lens, int64_t max_query_len,
                       const std::optional<torch::Tensor>& alibi_slopes,
                       const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                       torch::Tensor& v_scale, int64_t heads_per_tile) {
  check_cuda(query, "query");
  const int kv = kv_code(kv_cache_dtype);
  const int st = scalar_code(query.scalar_type());
  check_cache_dtype(key_cache, query, kv, "key_cache");
  check_cache_dtype(value_cache, query, kv, "value_cache");
  for (const torch::Tensor* t : {&out, &key_cache, &value_cache, &block_tables, &query_start_loc,
                                 &seq_lens, &k_scale, &v_scale})
    TORCH_CHECK(t->is_cuda() && t->get_device() == query.get_device(),
                "prefill_attention: every tensor must be on the query's CUDA device");
  TORCH_CHECK(!alibi_slopes || (alibi_slopes->is_cuda() &&
                                alibi_slopes->get_device() == query.get_device() &&
                                alibi_slopes->scalar_type() == at::ScalarType::Float),
              "alibi_slopes must be a float32 tensor on the query's device");
  TORCH_CHECK(k_scale.scalar_type() == at::ScalarType::Float && k_scale.numel() >= 1 &&
                  v_scale.scalar_type() == at::ScalarType::Float && v_scale.numel() >= 1,
              "k_scale / v_scale must be float32 tensors");
  TORCH_CHECK(query.dim() == 3 && out.dim() == 3, "query / out must be [num_tokens, num_heads, head_size]");
  TORCH_CHECK(out.scalar_type() == query.scalar_type() && out.sizes() == query.sizes(),
              "out must have query's type and shape");
  TORCH_CHECK(key_cache.dim() == 5 && value_cache.dim() == 4,
              "key_cache [blocks, kv_heads, head_size/x, block_size, x], value_cache [blocks, "
              "kv_heads, head_size, block_size]");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int, "block_tables must be int32");
  TORCH_CHECK(seq_lens.scalar_type() == at::ScalarType::Int, "seq_lens must be int32");
  TORCH_CHECK(query_start_loc.scalar_type() == at::ScalarType::Int, "query_start_loc must be int32");
  TORCH_CHECK(query_start_loc.is_contiguous() && seq_lens.is_contiguous() &&
              block_tables.stride(1) == 1, "metadata tensors must be contiguous");
  TORCH_CHECK(query_start_loc.numel() == seq_lens.numel() + 1,
              "query_start_loc must have num_seqs + 1 entries");
  TORCH_CHECK(block_tables.size(0) >= seq_lens.numel(), "block_tables has fewer rows than seq_lens");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == query.size(2),
              "query heads must be densely packed");
  TORCH_CHECK(out.stride(2) == 1 && out.stride(1) == out.size(2), "out heads must be densely packed");
  float* lse_ptr = nullptr;
  if (lse) {
    TORCH_CHECK(lse->scalar_type() == at::ScalarType::Float && lse->is_contiguous() &&
                lse->dim() == 2 && lse->size(0) == query.size(0) && lse->size(1) == query.size(1),
                "lse must be a contiguous float32 [num_tokens, num_heads]");
    lse_ptr = lse->data_ptr<float>();
  }
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  const float* alibi = alibi_slopes ? reinterpret_cast<const float*>(alibi_slopes->data_ptr()) : nullptr;
  check_status(b200pa_prefill_attention(
      out.data_ptr(), lse_ptr, query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
      (int)seq_lens.numel(), (int)query.size(1), (int)key_cache.size(1), (int)query.size(2),
      (int)key_cache.size(3), (float)scale, block_tables.data_ptr<int>(),
      (int)block_tables.stride(0), query_start_loc.data_ptr<int>(), seq_lens.data_ptr<int>(),
      (int)max_query_len, alibi, query.stride(0), out.stride(0), key_cache.stride(0),
      key_cache.stride(1), st, kv, reinterpret_cast<const float*>(k_scale.data_ptr()),
      reinterpret_cast<const float*>(v_scale.data_ptr()), (int)heads_per_tile, stream));
}

void merge_attn_states(torch::Tensor& output, std::optional<torch::Tensor> output_lse,
                       const torch::Tensor& prefix_output, const torch::Tensor& prefix_lse,
                       const torch::Tensor& suffix_output, const torch::Tensor& suffix_lse) {
  check_cuda(output, "output");
  const at::ScalarType t = output.scalar_type();
  TORCH_CHECK(t == at::ScalarType::Float || t == at::ScalarType::Half ||
                  t == at::ScalarType::BFloat16,
              "Unsupported data type of O: ", t);
  const int st = scalar_code(t);
  const int64_t head_size = output.size(2);
  const int64_t pack_size = 16 / output.element_size();
  TORCH_CHECK(head_size % pack_size == 0, "headsize must be multiple of pack_size:", pack_size);
  TORCH_CHECK(output.stride(-2) == head_size && output.stride(-1) == 1,
              "output heads must be contiguous in memory");
  TORCH_CHECK(prefix_output.stride(-2) == head_size && prefix_output.stride(-1) == 1,
              "prefix_output heads must be contiguous in memory");
  TORCH_CHECK(suffix_output.stride(-2) == head_size && suffix_output.stride(-1) == 1,
              "suffix_output heads must be contiguous in memory");
  // the kernel indexes [token, head] rows densely, as the reference kernel does (:37-38)
  TORCH_CHECK(output.is_contiguous() && prefix_output.is_contiguous() &&
                  suffix_output.is_contiguous() && prefix_lse.is_contiguous() &&
                  suffix_lse.is_contiguous(),
              "merge_attn_states works on contiguous tensors");
  TORCH_CHECK(prefix_output.scalar_type() == t && suffix_output.scalar_type() == t,
              "prefix_output / suffix_output dtype must match output");
  TORCH_CHECK(prefix_lse.scalar_type() == at::ScalarType::Float &&
                  suffix_lse.scalar_type() == at::ScalarType::Float,
              "prefix_lse / suffix_lse must be float32");
  float* out_lse = nullptr;
  if (output_lse.has_value()) {
    TORCH_CHECK(output_lse->scalar_type() == at::ScalarType::Float && output_lse->is_contiguous(),
                "output_lse must be contiguous float32");
    out_lse = output_lse->data_ptr<float>();
  }
  const at::cuda::OptionalCUDAGuard device_guard(device_of(prefix_output));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  check_status(b200pa_merge_attn_states(
      output.data_ptr(), out_lse, prefix_output.data_ptr(), prefix_lse.data_ptr<float>(),
      suffix_output.data_ptr(), suffix_lse.data_ptr<float>(), output.size(0), (int)output.size(1),
      (int)head_size, st, stream));
}

// vllm/attention/ops/common.py:86-116 of the reference: out [B,H,D] corrected IN PLACE with the
// all-gathered lses [N,B,H]; returns the total lse (empty_like(lses[0])).
torch::Tensor correct_attn_out(torch::Tensor& out, const torch::Tensor& lses, int64_t cp_rank) {
  check_cuda(out, "out");
  TORCH_CHECK(out.dim() == 3 && lses.dim() == 3 && lses.size(1) == out.size(0) &&
                  lses.size(2) == out.size(1),
              "correct_attn_out: out [B,H,D], lses [N,B,H]");
  TORCH_CHECK(lses.scalar_type() == at::ScalarType::Float, "lses must be float32");
  TORCH_CHECK(out.stride(2) == 1, "out head elements must be contiguous");
  TORCH_CHECK(cp_rank >= 0 && cp_rank < lses.size(0), "cp_rank out of range");
  const int st = scalar_code(out.scalar_type());
  const at::cuda::OptionalCUDAGuard device_guard(device_of(out));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  torch::Tensor lse = at::empty_strided({lses.size(1), lses.size(2)},
                                        {lses.stride(1), lses.stride(2)}, lses.options());
  check_status(b200pa_correct_attn_out(
      out.data_ptr(), lse.data_ptr<float>(), lses.data_ptr<float>(), (int)lses.size(0),
      (int)cp_rank, (int)out.size(0), (int)out.size(1), (int)out.size(2), out.stride(0),
      out.stride(1), lses.stride(0), lses.stride(1), lses.stride(2), st, stream));
  return lse;
}

void reshape_and_cache(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache,
                       torch::Tensor& value_cache, torch::Tensor& slot_mapping,
                       const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                       torch::Tensor& v_scale) {
  check_cuda(key, "key");
  const int kv = kv_code(kv_cache_dtype);
  const int st = scalar_code(key.scalar_type());
  check_cache_dtype(key_cache, key, kv, "key_cache");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long, "slot_mapping must be int64");
  TORCH_CHECK(key_cache.is_contiguous() && value_cache.is_contiguous(),
              "reshape_and_cache indexes the caches densely (as the reference kernel does)");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(key));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  check_status(b200pa_reshape_and_cache(
      key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
      slot_mapping.data_ptr<int64_t>(), (int)slot_mapping.size(0), key.stride(0), value.stride(0),
      (int)key.size(1), (int)key.size(2), (int)key_cache.size(3), (int)key_cache.size(4), st, kv,
      reinterpret_cast<const float*>(k_scale.data_ptr()),
      reinterpret_cast<const float*>(v_scale.data_ptr()), stream));
}

void reshape_and_cache_flash(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache,
                             torch::Tensor& value_cache, torch::Tensor& slot_mapping,
                             const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                             torch::Tensor& v_scale) {
  check_cuda(key, "key");
  const int kv = kv_code(kv_cache_dtype);
  const int st = scalar_code(key.scalar_type());
  check_cache_dtype(key_cache, key, kv, "key_cache");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long, "slot_mapping must be int64");
  TORCH_CHECK(key_cache.stride(0) == value_cache.stride(0));  // cache_kernels.cu:488
  TORCH_CHECK(key_cache.stride(3) == 1, "head elements must be contiguous in the cache");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(key));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  check_status(b200pa_reshape_and_cache_flash(
      key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
      slot_mapping.data_ptr<int64_t>(), (int)slot_mapping.size(0), key.stride(0), value.stride(0),
      key_cache.stride(0), key_cache.stride(1), key_cache.stride(2), (int)key.size(1),
      (int)key.size(2), (int)key_cache.size(1), st, kv,
      reinterpret_cast<const float*>(k_scale.data_ptr()),
      reinterpret_cast<const float*>(v_scale.data_ptr()), stream));
}

void convert_fp8(torch::Tensor& dst_cache, torch::Tensor& src_cache, const double scale,
                 const std::string& kv_cache_dtype) {
  TORCH_CHECK(src_cache.is_cuda(), "src must be on a GPU");
  TORCH_CHECK(dst_cache.is_cuda(), "dst must be on a GPU");
  TORCH_CHECK(src_cache.device().index() == dst_cache.device().index(),
              "src and dst must be on the same GPU");
  TORCH_CHECK(src_cache.is_contiguous() && dst_cache.is_contiguous(),
              "convert_fp8 works on contiguous caches");
  TORCH_CHECK(src_cache.numel() == dst_cache.numel(), "src and dst must have the same size");
  // the reference only ever instantiates the fp8 conversions (its "auto" branch
  // reaches an assert(false) in scaled_convert, quant_utils.cuh:528-539)
  TORCH_CHECK(kv_cache_dtype == "fp8" || kv_cache_dtype == "fp8_e4m3" || kv_cache_dtype == "fp8_e5m2",
              "Unsupported data type: ", kv_cache_dtype);
  const int kv = kv_code(kv_cache_dtype);
  const bool to_fp8 = dst_cache.element_size() == 1;
  TORCH_CHECK(to_fp8 || src_cache.element_size() == 1, "one side of convert_fp8 must be uint8");
  const int st = scalar_code(to_fp8 ? src_cache.scalar_type() : dst_cache.scalar_type());
  at::cuda::OptionalCUDAGuard device_guard(src_cache.device());
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  check_status(b200pa_convert_fp8(dst_cache.data_ptr(), src_cache.data_ptr(), src_cache.numel(),
                                  (float)scale, st, kv, to_fp8 ? 1 : 0, stream));
}

// csrc/cache_kernels.cu:27-70 of the reference: same device rules, same messages; block_mapping
// must be a CPU tensor [num_pairs, 2] int64.
void swap_blocks(torch::Tensor& src, torch::Tensor& dst, const torch::Tensor& block_mapping) {
  const torch::Device sd = src.device(), dd = dst.device();
  int direction;
  if (sd.is_cuda() && dd.is_cuda()) {
    TORCH_CHECK(sd.index() == dd.index(), "src and dst must be on the same GPU");
    direction = B200PA_COPY_D2D;
  } else if (sd.is_cuda() && dd.is_cpu()) {
    direction = B200PA_COPY_D2H;
  } else if (sd.is_cpu() && dd.is_cuda()) {
    direction = B200PA_COPY_H2D;
  } else {
    TORCH_CHECK(false, "Invalid device combination");
  }
  TORCH_CHECK(block_mapping.device().is_cpu(), "block_mapping must be on CPU");
  TORCH_CHECK(block_mapping.scalar_type() == at::ScalarType::Long && block_mapping.dim() == 2 &&
                  block_mapping.size(1) == 2,
              "block_mapping must be int64 [num_pairs, 2]");
  const torch::Tensor map = block_mapping.contiguous();
  const int64_t block_bytes = src.element_size() * src.stride(0);
  const at::cuda::OptionalCUDAGuard device_guard(sd.is_cuda() ? sd : dd);
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  check_status(b200pa_swap_blocks(src.data_ptr(), dst.data_ptr(), block_bytes,
                                  map.data_ptr<int64_t>(), map.size(0), direction, stream));
}

// csrc/cache_kernels.cu:125-172 of the reference (block_mapping on the device, [num_pairs, 2])
void copy_blocks(std::vector<torch::Tensor> const& key_caches,
                 std::vector<torch::Tensor> const& value_caches,
                 const torch::Tensor& block_mapping) {
  const int num_layers = (int)key_caches.size();
  TORCH_CHECK(num_layers == (int)value_caches.size());
  if (num_layers == 0) return;
  const torch::Device dev = key_caches[0].device();
  TORCH_CHECK(dev.is_cuda());
  TORCH_CHECK(block_mapping.is_cuda() && block_mapping.scalar_type() == at::ScalarType::Long &&
                  block_mapping.is_contiguous(),
              "block_mapping must be a contiguous int64 tensor on the GPU");
  std::vector<void*> kp(num_layers), vp(num_layers);
  for (int l = 0; l < num_layers; ++l) {
    kp[l] = key_caches[l].data_ptr();
    vp[l] = value_caches[l].data_ptr();
  }
  const int64_t block_bytes = key_caches[0][0].numel() * key_caches[0].element_size();
  const at::cuda::OptionalCUDAGuard device_guard(dev);
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  check_status(b200pa_copy_blocks(kp.data(), vp.data(), num_layers,
                                  block_mapping.data_ptr<int64_t>(), (int)block_mapping.size(0),
                                  block_bytes, stream));
}

static void check_gather_tensors(torch::Tensor const& src_cache, torch::Tensor const& dst,
                                 torch::Tensor const& block_table,
                                 torch::Tensor const& cu_seq_lens,
                                 const std::optional<torch::Tensor>& seq_starts) {
  // csrc/cache_kernels.cu:748-768 / :879-899 of the reference
  TORCH_CHECK(block_table.dtype() == torch::kInt32, "block_table must be int32");
  TORCH_CHECK(cu_seq_lens.dtype() == torch::kInt32, "cu_seq_lens must be int32");
  if (seq_starts.has_value()) {
    TORCH_CHECK(seq_starts.value().dtype() == torch::kInt32, "seq_starts must be int32");
  }
  TORCH_CHECK(src_cache.device() == dst.device(), "src_cache and dst must be on the same device");
  TORCH_CHECK(src_cache.device() == block_table.device(),
              "src_cache and block_table must be on the same device");
  TORCH_CHECK(src_cache.device() == cu_seq_lens.device(),
              "src_cache and cu_seq_lens must be on the same device");
  if (seq_starts.has_value()) {
    TORCH_CHECK(src_cache.device() == seq_starts.value().device(),
                "src_cache and seq_starts must be on the same device");
  }
  TORCH_CHECK(src_cache.is_cuda(), "src_cache must be on a GPU (this package has no CPU path)");
  TORCH_CHECK(block_table.stride(-1) == 1 && cu_seq_lens.is_contiguous(),
              "block_table rows and cu_seq_lens must be dense");
}

void gather_and_maybe_dequant_cache(torch::Tensor const& src_cache, torch::Tensor const& dst,
                                    torch::Tensor const& block_table,
                                    torch::Tensor const& cu_seq_lens, int64_t batch_size,
                                    const std::string& kv_cache_dtype, torch::Tensor const& scale,
                                    std::optional<torch::Tensor> seq_starts) {
  check_gather_tensors(src_cache, dst, block_table, cu_seq_lens, seq_starts);
  const int kv = kv_code(kv_cache_dtype);
  const int st = scalar_code(dst.scalar_type());
  if (kv == B200PA_KV_AUTO) {
    TORCH_CHECK(src_cache.scalar_type() == dst.scalar_type(),
                "kv_cache_dtype \"auto\": src_cache and dst must have the same dtype");
  } else {
    TORCH_CHECK(src_cache.element_size() == 1, "fp8 kv cache must be a 1-byte tensor");
  }
  const int64_t entry_size = src_cache.flatten(2, -1).size(2);
  TORCH_CHECK(src_cache.dim() == 3 || src_cache.flatten(2, -1).stride(2) == 1,
              "cache entries must be dense");
  at::cuda::OptionalCUDAGuard device_guard(src_cache.device());
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  check_status(b200pa_gather_and_maybe_dequant_cache(
      src_cache.data_ptr(), dst.data_ptr(), block_table.data_ptr<int32_t>(),
      cu_seq_lens.data_ptr<int32_t>(), (int)batch_size, (int)src_cache.size(1), (int)entry_size,
      block_table.stride(0), src_cache.stride(0), src_cache.stride(1), dst.stride(0), st, kv,
      reinterpret_cast<const float*>(scale.data_ptr()),
      seq_starts.has_value() ? seq_starts.value().data_ptr<int32_t>() : nullptr, dst.size(0),
      stream));
}

void cp_gather_cache(torch::Tensor const& src_cache, torch::Tensor const& dst,
                     torch::Tensor const& block_table, torch::Tensor const& cu_seq_lens,
                     int64_t batch_size, std::optional<torch::Tensor> seq_starts) {
  check_gather_tensors(src_cache, dst, block_table, cu_seq_lens, seq_starts);
  TORCH_CHECK(src_cache.dtype() == dst.dtype(), "src_cache and dst must have the same dtype");
  const int eb = (int)src_cache.element_size();
  TORCH_CHECK(eb == 1 || eb == 2 || eb == 4, "Unsupported data type width: ", eb * 8);
  const int64_t entry_size = src_cache.flatten(2, -1).size(2);
  at::cuda::OptionalCUDAGuard device_guard(src_cache.device());
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  check_status(b200pa_cp_gather_cache(
      src_cache.data_ptr(), dst.data_ptr(), block_table.data_ptr<int32_t>(),
      cu_seq_lens.data_ptr<int32_t>(), (int)batch_size, (int)src_cache.size(1), (int)entry_size,
      block_table.stride(0), src_cache.stride(0), src_cache.stride(1), dst.stride(0), eb,
      seq_starts.has_value() ? seq_starts.value().data_ptr<int32_t>() : nullptr, dst.size(0),
      stream));
}

int64_t get_max_shared_memory_per_block_device_attribute(int64_t device_id) {
  int64_t v = 0;
  check_status(b200pa_get_max_shared_memory_per_block_device_attribute((int)device_id, &v));
  return v;
}

constexpr const char* kSchemaV1 =
    "paged_attention_v1("
    "    Tensor! out, Tensor query, Tensor key_cache,"
    "    Tensor value_cache, int num_kv_heads, float scale,"
    "    Tensor block_tables, Tensor seq_lens, int block_size,"
    "    int max_seq_len, Tensor? alibi_slopes,"
    "    str kv_cache_dtype, Tensor k_scale, Tensor v_scale,"
    "    int tp_rank, int blocksparse_local_blocks,"
    "    int blocksparse_vert_stride, int blocksparse_block_size,"
    "    int blocksparse_head_sliding_step) -> ()";
constexpr const char* kSchemaV2 =
    "paged_attention_v2("
    "    Tensor! out, Tensor! exp_sums, Tensor! max_logits,"
    "    Tensor! tmp_out, Tensor query, Tensor key_cache,"
    "    Tensor value_cache, int num_kv_heads, float scale,"
    "    Tensor block_tables, Tensor seq_lens, int block_size,"
    "    int max_seq_len, Tensor? alibi_slopes,"
    "    str kv_cache_dtype, Tensor k_scale, Tensor v_scale,"
    "    int tp_rank, int blocksparse_local_blocks,"
    "    int blocksparse_vert_stride, int blocksparse_block_size,"
    "    int blocksparse_head_sliding_step) -> ()";
constexpr const char* kSchemaMerge =
    "merge_attn_states("
    "    Tensor! output,"
    "    Tensor!? output_lse,"
    "    Tensor prefix_output,"
    "    Tensor prefix_lse,"
    "    Tensor suffix_output,"
    "    Tensor suffix_lse) -> ()";
constexpr const char* kSchemaLse =
    "paged_attention_lse("
    "    Tensor! out, Tensor! lse, Tensor query, Tensor key_cache,"
    "    Tensor value_cache, int num_kv_heads, float scale,"
    "    Tensor block_tables, Tensor seq_lens, int block_size,"
    "    int max_seq_len, Tensor? alibi_slopes,"
    "    str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()";
constexpr const char* kSchemaStep =
    "decode_step("
    "    Tensor! out, Tensor!? lse, Tensor query, Tensor key, Tensor value,"
    "    Tensor! key_cache, Tensor! value_cache, Tensor slot_mapping,"
    "    int num_kv_heads, float scale,"
    "    Tensor block_tables, Tensor seq_lens, int block_size,"
    "    int max_seq_len, Tensor? alibi_slopes,"
    "    str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()";
constexpr const char* kSchemaFlash =
    "paged_attention_flash("
    "    Tensor! out, Tensor!? lse, Tensor query, Tensor? key, Tensor? value,"
    "    Tensor! key_cache, Tensor! value_cache, Tensor? slot_mapping, float scale,"
    "    Tensor block_tables, Tensor seq_lens, int max_seq_len, Tensor? alibi_slopes,"
    "    str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()";
constexpr const char* kSchemaPrefill =
    "prefill_attention("
    "    Tensor! out, Tensor!? lse, Tensor query, Tensor key_cache, Tensor value_cache,"
    "    float scale, Tensor block_tables, Tensor query_start_loc, Tensor seq_lens,"
    "    int max_query_len, Tensor? alibi_slopes, str kv_cache_dtype,"
    "    Tensor k_scale, Tensor v_scale, int heads_per_tile=0) -> ()";
constexpr const char* kSchemaCorrect =
    "correct_attn_out(Tensor! out, Tensor lses, int cp_rank) -> Tensor";
constexpr const char* kSchemaReshape =
    "reshape_and_cache(Tensor key, Tensor value,"
    "                  Tensor! key_cache, Tensor! value_cache,"
    "                  Tensor slot_mapping,"
    "                  str kv_cache_dtype,"
    "                  Tensor k_scale, Tensor v_scale) -> ()";
constexpr const char* kSchemaReshapeFlash =
    "reshape_and_cache_flash(Tensor key, Tensor value,"
    "                        Tensor! key_cache,"
    "                        Tensor! value_cache,"
    "                        Tensor slot_mapping,"
    "                        str kv_cache_dtype,"
    "                        Tensor k_scale, Tensor v_scale) -> ()";
constexpr const char* kSchemaConvert =
    "convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale, str kv_cache_dtype) -> ()";
constexpr const char* kSchemaSmem =
    "get_max_shared_memory_per_block_device_attribute(int device_id) -> int";

void define_attention(torch::Library& ops) {
  ops.def(kSchemaV1);
  ops.impl("paged_attention_v1", torch::kCUDA, &paged_attention_v1);
  ops.def(kSchemaV2);
  ops.impl("paged_attention_v2", torch::kCUDA, &paged_attention_v2);
  ops.def(kSchemaMerge);
  ops.impl("merge_attn_states", torch::kCUDA, &merge_attn_states);
}
void define_extras(torch::Library& ops) {
  {
    std::string schema(kSchemaV2);
    schema.replace(0, std::string("paged_attention_v2").size(), "paged_attention_v2_partials");
    ops.def(schema.c_str());
    ops.impl("paged_attention_v2_partials", torch::kCUDA, &paged_attention_v2_partials);
  }
  ops.def(kSchemaLse);
  ops.impl("paged_attention_lse", torch::kCUDA, &paged_attention_lse);
  ops.def(kSchemaStep);
  ops.impl("decode_step", torch::kCUDA, &decode_step);
  ops.def(kSchemaFlash);
  ops.impl("paged_attention_flash", torch::kCUDA, &paged_attention_flash);
  ops.def(kSchemaPrefill);
  ops.impl("prefill_attention", torch::kCUDA, &prefill_attention);
  ops.def(kSchemaCorrect);
  ops.impl("correct_attn_out", torch::kCUDA, &correct_attn_out);
}
void define_cache(torch::Library& ops) {
  // csrc/torch_bindings.cpp:652-660, 703-715 of the reference
  ops.def("swap_blocks(Tensor src, Tensor! dst, Tensor block_mapping) -> ()");
  ops.impl("swap_blocks", torch::kCUDA, &swap_blocks);
  ops.def("copy_blocks(Tensor(a!)[] key_caches, Tensor[](b!) value_caches, "
          "Tensor block_mapping) -> ()");
  ops.impl("copy_blocks", torch::kCUDA, &copy_blocks);
  ops.def("gather_and_maybe_dequant_cache(Tensor src_cache, Tensor! dst, "
          "                               Tensor block_table, Tensor cu_seq_lens, "
          "                               int batch_size, "
          "                               str kv_cache_dtype, "
          "                               Tensor scale, Tensor? seq_starts) -> ()");
  ops.impl("gather_and_maybe_dequant_cache", torch::kCUDA, &gather_and_maybe_dequant_cache);
  ops.def("cp_gather_cache(Tensor src_cache, Tensor! dst, Tensor block_table, "
          "Tensor cu_seq_lens, int batch_size, Tensor? seq_starts) -> ()");
  ops.impl("cp_gather_cache", torch::kCUDA, &cp_gather_cache);
  ops.def(kSchemaReshape);
  ops.impl("reshape_and_cache", torch::kCUDA, &reshape_and_cache);
  ops.def(kSchemaReshapeFlash);
  ops.impl("reshape_and_cache_flash", torch::kCUDA, &reshape_and_cache_flash);
  ops.def(kSchemaConvert);
  ops.impl("convert_fp8", torch::kCUDA, &convert_fp8);
}
void define_utils(torch::Library& ops) {
  ops.def(kSchemaSmem);
  // no tensor argument -> no dispatch key to infer: register as a catch-all like the reference
  ops.impl("get_max_shared_memory_per_block_device_attribute",
           &get_max_shared_memory_per_block_device_attribute);
}

std::vector<std::unique_ptr<torch::Library>> g_vllm_libs;

}  // namespace

TORCH_LIBRARY(b200pa, ops) {
  define_attention(ops);
  define_extras(ops);
  define_cache(ops);
  define_utils(ops);
}

// Register the same ops under the reference's namespaces.  Returns 0 on success,
// 1 if they were already registered by this call earlier.  Throws (c10::Error)
// if another extension already defined the schemas in this process.
extern "C" int b200pa_register_vllm_namespaces(void) {
  if (!g_vllm_libs.empty()) return 1;
  auto mk = [](const char* ns) {
    return std::make_unique<torch::Library>(torch::Library::FRAGMENT, ns, std::nullopt, __FILE__,
                                            __LINE__);
  };
  auto c = mk("_C");
  define_attention(*c);
  auto cache = mk("_C_cache_ops");
  define_cache(*cache);
  auto utils = mk("_C_cuda_utils");
  define_utils(*utils);
  g_vllm_libs.push_back(std::move(c));
  g_vllm_libs.push_back(std::move(cache));
  g_vllm_libs.push_back(std::move(utils));
  return 0;
}
