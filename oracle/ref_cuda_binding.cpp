// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// Op registration for the REFERENCE's own CUDA kernels of the paged-decode
// path, compiled unmodified from where they lie under /root/reference/csrc
// (attention/paged_attention_v1.cu, attention/paged_attention_v2.cu,
// cache_kernels.cu; recipe: oracle/build_ref_cuda.sh).  This is oracle tier O1
// of SURVEY.md section 8c: the parity target named by north_star ("match the
// reference's own paged_attention") and the same-box "kernel to beat".  Only
// tests/ and tools/ (comparison timing) load it; the product never does.
//
// Schemas restate csrc/torch_bindings.cpp:52-76 and :667-716 of the reference,
// under the private namespace `vllm_ref_cuda` so they can never shadow `_C`.
#include <torch/all.h>
#include <torch/library.h>

#include <optional>
#include <string>
#include <vector>

// csrc/attention/paged_attention_v1.cu:160, paged_attention_v2.cu:167
void paged_attention_v1(
    torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache,
    torch::Tensor& value_cache, int64_t num_kv_heads, double scale,
    torch::Tensor& block_tables, torch::Tensor& seq_lens, int64_t block_size,
    int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
    const std::string& kv_cache_dtype, torch::Tensor& k_scale,
    torch::Tensor& v_scale, const int64_t tp_rank,
    const int64_t blocksparse_local_blocks,
    const int64_t blocksparse_vert_stride, const int64_t blocksparse_block_size,
    const int64_t blocksparse_head_sliding_step);

void paged_attention_v2(
    torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits,
    torch::Tensor& tmp_out, torch::Tensor& query, torch::Tensor& key_cache,
    torch::Tensor& value_cache, int64_t num_kv_heads, double scale,
    torch::Tensor& block_tables, torch::Tensor& seq_lens, int64_t block_size,
    int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
    const std::string& kv_cache_dtype, torch::Tensor& k_scale,
    torch::Tensor& v_scale, const int64_t tp_rank,
    const int64_t blocksparse_local_blocks,
    const int64_t blocksparse_vert_stride, const int64_t blocksparse_block_size,
    const int64_t blocksparse_head_sliding_step);

// csrc/cache_kernels.cu:416, :459, :573
void reshape_and_cache(torch::Tensor& key, torch::Tensor& value,
                       torch::Tensor& key_cache, torch::Tensor& value_cache,
                       torch::Tensor& slot_mapping,
                       const std::string& kv_cache_dtype,
                       torch::Tensor& k_scale, torch::Tensor& v_scale);
void reshape_and_cache_flash(torch::Tensor& key, torch::Tensor& value,
                             torch::Tensor& key_cache,
                             torch::Tensor& value_cache,
                             torch::Tensor& slot_mapping,
                             const std::string& kv_cache_dtype,
                             torch::Tensor& k_scale, torch::Tensor& v_scale);
void convert_fp8(torch::Tensor& dst_cache, torch::Tensor& src_cache,
                 const double scale, const std::string& kv_cache_dtype);

// csrc/attention/merge_attn_states.cu:172 (declared csrc/ops.h:56-61)
void merge_attn_states(torch::Tensor& output,
                       std::optional<torch::Tensor> output_lse,
                       const torch::Tensor& prefix_output,
                       const torch::Tensor& prefix_lse,
                       const torch::Tensor& suffix_output,
                       const torch::Tensor& suffix_lse);

// csrc/cache_kernels.cu:27, :125, :734, :865 (declared csrc/cache.h:8-60)
void swap_blocks(torch::Tensor& src, torch::Tensor& dst,
                 const torch::Tensor& block_mapping);
void copy_blocks(std::vector<torch::Tensor> const& key_caches,
                 std::vector<torch::Tensor> const& value_caches,
                 const torch::Tensor& block_mapping);
void gather_and_maybe_dequant_cache(
    torch::Tensor const& src_cache, torch::Tensor const& dst,
    torch::Tensor const& block_table, torch::Tensor const& cu_seq_lens,
    int64_t batch_size, const std::string& kv_cache_dtype,
    torch::Tensor const& scale, std::optional<torch::Tensor> seq_starts);
void cp_gather_cache(torch::Tensor const& src_cache, torch::Tensor const& dst,
                     torch::Tensor const& block_table,
                     torch::Tensor const& cu_seq_lens, int64_t batch_size,
                     std::optional<torch::Tensor> seq_starts);

#define ATTN_TAIL                                                        \
  "Tensor query, Tensor key_cache, Tensor value_cache, int num_kv_heads," \
  " float scale, Tensor block_tables, Tensor seq_lens, int block_size,"  \
  " int max_seq_len, Tensor? alibi_slopes, str kv_cache_dtype,"          \
  " Tensor k_scale, Tensor v_scale, int tp_rank,"                        \
  " int blocksparse_local_blocks, int blocksparse_vert_stride,"          \
  " int blocksparse_block_size, int blocksparse_head_sliding_step) -> ()"

TORCH_LIBRARY(vllm_ref_cuda, ops) {
  ops.def("paged_attention_v1(Tensor! out, " ATTN_TAIL);
  ops.impl("paged_attention_v1", torch::kCUDA, &paged_attention_v1);
  ops.def(
      "paged_attention_v2(Tensor! out, Tensor! exp_sums, Tensor! max_logits,"
      " Tensor! tmp_out, " ATTN_TAIL);
  ops.impl("paged_attention_v2", torch::kCUDA, &paged_attention_v2);
  ops.def(
      "reshape_and_cache(Tensor key, Tensor value, Tensor! key_cache,"
      " Tensor! value_cache, Tensor slot_mapping, str kv_cache_dtype,"
      " Tensor k_scale, Tensor v_scale) -> ()");
  ops.impl("reshape_and_cache", torch::kCUDA, &reshape_and_cache);
  ops.def(
      "reshape_and_cache_flash(Tensor key, Tensor value, Tensor! key_cache,"
      " Tensor! value_cache, Tensor slot_mapping, str kv_cache_dtype,"
      " Tensor k_scale, Tensor v_scale) -> ()");
  ops.impl("reshape_and_cache_flash", torch::kCUDA, &reshape_and_cache_flash);
  ops.def(
      "convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale,"
      " str kv_cache_dtype) -> ()");
  ops.impl("convert_fp8", torch::kCUDA, &convert_fp8);
  ops.def(
      "merge_attn_states(Tensor! output, Tensor!? output_lse,"
      " Tensor prefix_output, Tensor prefix_lse, Tensor suffix_output,"
      " Tensor suffix_lse) -> ()");
  ops.impl("merge_attn_states", torch::kCUDA, &merge_attn_states);
  ops.def("swap_blocks(Tensor src, Tensor! dst, Tensor block_mapping) -> ()");
  ops.impl("swap_blocks", torch::kCUDA, &swap_blocks);
  ops.def(
      "copy_blocks(Tensor(a!)[] key_caches, Tensor[](b!) value_caches,"
      " Tensor block_mapping) -> ()");
  ops.impl("copy_blocks", torch::kCUDA, &copy_blocks);
  ops.def(
      "gather_and_maybe_dequant_cache(Tensor src_cache, Tensor! dst,"
      " Tensor block_table, Tensor cu_seq_lens, int batch_size,"
      " str kv_cache_dtype, Tensor scale, Tensor? seq_starts) -> ()");
  ops.impl("gather_and_maybe_dequant_cache", torch::kCUDA,
           &gather_and_maybe_dequant_cache);
  ops.def(
      "cp_gather_cache(Tensor src_cache, Tensor! dst, Tensor block_table,"
      " Tensor cu_seq_lens, int batch_size, Tensor? seq_starts) -> ()");
  ops.impl("cp_gather_cache", torch::kCUDA, &cp_gather_cache);
}
