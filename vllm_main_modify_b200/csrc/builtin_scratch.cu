// b200pa_paged_attention_v1_builtin_scratch: paged_attention_v1 for callers that have no way to
// hand this library a workspace.
//
// The decode kernel behind b200pa_paged_attention_v1 keeps the partial results of sequences cut by
// CTA boundaries, and the ticket counters that elect the merging CTA, in a `workspace` the caller
// allocates (about 23.5 MiB on a B200 for head size 128).  Not every host of this ABI can do that:
//   - a plugin loaded into an engine that has already given all free HBM to its KV-cache pool
//     (the reference sizes the cache from whatever is free after profiling, so a cudaMalloc made
//     afterwards can simply fail) and owns every byte through its own arena;
//   - an FFI host (cgo, JNI, a ctypes script) that is handed device pointers by someone else and
//     has no CUDA runtime binding of its own to allocate with;
//   - a latency-critical decode loop that must never hit cudaMalloc's implicit device-wide
//     synchronisation.
// For them the workspace lives HERE, in a file-scope device array that is part of the library
// image: every GPU that loads libb200pa.so holds B200PA_BUILTIN_SCRATCH_BYTES of it from module
// load until the process exits, whether or not this entry point is ever called.
//
// This file never calls cudaMalloc / cudaFree or anything that allocates; it launches the same
// kernels as b200pa_paged_attention_v1 (through that entry point), so results are the same.
#include <atomic>

#include "common.cuh"

// The scratch itself.  In device code this name is the array; in HOST code it is only a
// stand-in symbol of the host image, NOT a device address -- it must never be passed to a
// kernel.  Host code gets the real address from cudaGetSymbolAddress (scratch_of_current_device).
static __device__ __align__(256) unsigned char g_builtin_scratch[B200PA_BUILTIN_SCRATCH_BYTES];

namespace b200pa {
namespace {

// Device address of g_builtin_scratch, per device ordinal: every device has its own instance of
// the array (at its own address), so the look-up is made once per device with that device current.
constexpr int kMaxDevices = 64;
std::atomic<void*> g_scratch_addr[kMaxDevices];

int scratch_of_current_device(void** addr) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return set_error(B200PA_ERR_CUDA, "builtin scratch: cudaGetDevice: %s", cudaGetErrorString(e));
  }
  if (dev < 0 || dev >= kMaxDevices)
    return set_error(B200PA_ERR_UNSUPPORTED, "builtin scratch: device ordinal %d out of range", dev);
  void* p = g_scratch_addr[dev].load(std::memory_order_acquire);
  if (!p) {
    // set-up, once per device (two threads racing here both get the same answer)
    e = cudaGetSymbolAddress(&p, g_builtin_scratch);
    if (e != cudaSuccess || !p) {
      (void)cudaGetLastError();
      return set_error(B200PA_ERR_CUDA, "builtin scratch: cudaGetSymbolAddress on device %d: %s",
                       dev, cudaGetErrorString(e));
    }
    g_scratch_addr[dev].store(p, std::memory_order_release);
  }
  *addr = p;
  return B200PA_OK;
}

}  // namespace
}  // namespace b200pa

using namespace b200pa;

extern "C" int b200pa_paged_attention_v1_builtin_scratch(
    void* out, const void* query, const void* key_cache, const void* value_cache, int num_seqs,
    int num_heads, int num_kv_heads, int head_size, int block_size, float scale,
    const int32_t* block_tables, int max_num_blocks_per_seq, const int32_t* seq_lens,
    int max_seq_len, const float* alibi_slopes, int64_t q_stride, int64_t kv_block_stride,
    int64_t kv_head_stride, int scalar_type, int kv_dtype, const float* k_scale,
    const float* v_scale, int tp_rank, int blocksparse_local_blocks, int blocksparse_vert_stride,
    int blocksparse_block_size, int blocksparse_head_sliding_step, void* stream) {
  // The scratch the kernels want depends on the head size and the SM count only (not on the batch
  // or the context), so this is the one size check there is.
  const int64_t need = b200pa_paged_attention_workspace_bytes(head_size);
  if (need > (int64_t)B200PA_BUILTIN_SCRATCH_BYTES)
    return set_error(B200PA_ERR_UNSUPPORTED,
                     "paged_attention_v1_builtin_scratch: head size %d needs %lld bytes of scratch "
                     "on this device, the built-in scratch holds %lld (head size <= 128 on a B200); "
                     "use b200pa_paged_attention_v1 with a caller-allocated workspace",
                     head_size, (long long)need, (long long)B200PA_BUILTIN_SCRATCH_BYTES);
  void* scratch = nullptr;
  int rc = scratch_of_current_device(&scratch);
  if (rc) return rc;
  // The array is shared by every call and keeps what the previous one left.  Only the ticket
  // counters at its front must read zero when the kernel starts; clear them on the caller's
  // stream, ordered before the launch (the rest is written before it is read).
  const cudaError_t e = cudaMemsetAsync(scratch, 0, B200PA_WORKSPACE_HEADER_BYTES,
                                        reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return set_error(B200PA_ERR_CUDA, "builtin scratch: cudaMemsetAsync: %s", cudaGetErrorString(e));
  }
  return b200pa_paged_attention_v1(
      out, query, key_cache, value_cache, num_seqs, num_heads, num_kv_heads, head_size, block_size,
      scale, block_tables, max_num_blocks_per_seq, seq_lens, max_seq_len, alibi_slopes, q_stride,
      kv_block_stride, kv_head_stride, scalar_type, kv_dtype, k_scale, v_scale, tp_rank,
      blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
      blocksparse_head_sliding_step, scratch, (int64_t)B200PA_BUILTIN_SCRATCH_BYTES, stream);
}
