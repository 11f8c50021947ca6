// Block-management ops on the paged KV cache (sm_100a) -- SURVEY.md section 8f row 4.
//
//  * copy_blocks_kernel -- the job of the reference's copy_blocks (csrc/cache_kernels.cu:74-101,
//    launcher :125-172): for every layer and every (src, dst) pair copy one block of the key cache
//    and of the value cache.  The reference uploads the per-layer pointer arrays with a blocking
//    .to(device) (:148-153); here they travel as kernel parameters (chunks of kPtrChunk layers),
//    so the op neither allocates nor synchronises and can be captured in a CUDA graph.
//  * gather_cache_kernel -- the job of gather_and_maybe_dequant_cache (:628-712, launcher
//    :734-782) and cp_gather_cache (:785-849, launcher :865-919): gather the tokens of every
//    sequence from a [num_blocks, block_size, entry...] cache into a dense [tot_tokens, entry]
//    tensor through the block table, optionally dequantising fp8 -> scalar_t with the reference's
//    scaled_convert (quant_utils.cuh:362-428).  The reference cuts every sequence into a fixed
//    number of splits (grid (batch, 2|4|16)), so ragged batches leave CTAs idle; here all tokens
//    are flattened and dealt to warps grid-stride (token -> sequence by binary search in
//    cu_seq_lens), one warp per token entry, 16-byte accesses when the layout allows.
//
// All of it is HBM-bound byte moving: algorithmic bytes = bytes read + bytes written, no reuse.
#include "common.cuh"

namespace b200pa {

namespace {

constexpr int kCopyThreads = 256;
constexpr int kPtrChunk = 128;  // layers per launch: 2 x 128 pointers = 2 KiB of kernel parameters

struct LayerPtrs {
  void* key[kPtrChunk];
  void* value[kPtrChunk];
};

// grid (layers_in_chunk, num_pairs); block_bytes = numel_per_block * element size
template <typename V>
__global__ void __launch_bounds__(kCopyThreads)
copy_blocks_kernel(const LayerPtrs ptrs, const int64_t* __restrict__ block_mapping,
                   int64_t block_bytes) {
  const int layer = blockIdx.x, pair = blockIdx.y;
  const int64_t src = block_mapping[2 * pair], dst = block_mapping[2 * pair + 1];
  const int64_t n = block_bytes / (int64_t)sizeof(V);
  {
    const V* s = reinterpret_cast<const V*>(static_cast<const char*>(ptrs.key[layer]) + src * block_bytes);
    V* d = reinterpret_cast<V*>(static_cast<char*>(ptrs.key[layer]) + dst * block_bytes);
    for (int64_t i = threadIdx.x; i < n; i += kCopyThreads) d[i] = s[i];
  }
  {
    const V* s = reinterpret_cast<const V*>(static_cast<const char*>(ptrs.value[layer]) + src * block_bytes);
    V* d = reinterpret_cast<V*>(static_cast<char*>(ptrs.value[layer]) + dst * block_bytes);
    for (int64_t i = threadIdx.x; i < n; i += kCopyThreads) d[i] = s[i];
  }
}

// cache element / destination element conversions of the gather ops
template <int ST, int KV> struct GatherCvt {   // fp8 cache -> scalar_t, scaled (bit-faithful)
  using src_t = uint8_t;
  static __device__ __forceinline__ typename Scalar<ST>::type cvt(uint8_t v, float scale) {
    return Scalar<ST>::from_float(fp8_dequant_faithful<KV, ST>(v, scale));
  }
};
template <int ST> struct GatherCvt<ST, B200PA_KV_AUTO> {   // same type: plain copy
  using src_t = typename Scalar<ST>::type;
  static __device__ __forceinline__ typename Scalar<ST>::type cvt(src_t v, float) { return v; }
};

// Sequence of flattened token `tok`: largest b with cu_seq_lens[b] <= tok (empty sequences are
// skipped because their interval is empty).
__device__ __forceinline__ int seq_of_token(const int32_t* __restrict__ cu_seq_lens, int batch_size,
                                            int tok) {
  int lo = 0, hi = batch_size;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (cu_seq_lens[mid] <= tok) lo = mid; else hi = mid;
  }
  return lo;
}

// gather_and_maybe_dequant_cache: one warp per token; seq_starts / block_size whole blocks are
// skipped at the front of the block-table row (cache_kernels.cu:663-668 of the reference).
// VEC: source and destination entries are 16-byte aligned and entry_size fills whole vectors.
template <int ST, int KV, bool VEC>
__global__ void __launch_bounds__(kCopyThreads)
gather_cache_kernel(const typename GatherCvt<ST, KV>::src_t* __restrict__ src_cache,
                    typename Scalar<ST>::type* __restrict__ dst,
                    const int32_t* __restrict__ block_table, const int32_t* __restrict__ cu_seq_lens,
                    int batch_size, int block_size, int entry_size, int64_t block_table_stride,
                    int64_t cache_block_stride, int64_t cache_entry_stride,
                    int64_t dst_entry_stride, const float* __restrict__ scale,
                    const int32_t* __restrict__ seq_starts) {
  using S = typename GatherCvt<ST, KV>::src_t;
  using T = typename Scalar<ST>::type;
  const int lane = threadIdx.x & 31;
  const int warps = gridDim.x * (kCopyThreads / 32);
  const int total = cu_seq_lens[batch_size];
  float sc = 1.f;
  if constexpr (KV != B200PA_KV_AUTO) sc = *scale;
  for (int tok = blockIdx.x * (kCopyThreads / 32) + (threadIdx.x >> 5); tok < total; tok += warps) {
    const int b = seq_of_token(cu_seq_lens, batch_size, tok);
    const int pos = tok - cu_seq_lens[b];
    const int blk_off = seq_starts != nullptr ? seq_starts[b] / block_size : 0;
    const int page = pos / block_size, slot = pos - page * block_size;
    const int64_t block_id = block_table[(int64_t)b * block_table_stride + blk_off + page];
    const S* s = src_cache + block_id * cache_block_stride + (int64_t)slot * cache_entry_stride;
    T* d = dst + (int64_t)tok * dst_entry_stride;
    if constexpr (VEC) {
      if constexpr (KV == B200PA_KV_AUTO) {
        const int n = entry_size * (int)sizeof(T) / 16;
        for (int i = lane; i < n; i += 32)
          reinterpret_cast<uint4*>(d)[i] = reinterpret_cast<const uint4*>(s)[i];
      } else {
        // 16 fp8 codes in, 16 scalar_t out
        const int n = entry_size / 16;
        for (int i = lane; i < n; i += 32) {
          const uint4 raw = reinterpret_cast<const uint4*>(s)[i];
          const uint8_t* c = reinterpret_cast<const uint8_t*>(&raw);
          alignas(16) T o[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) o[j] = GatherCvt<ST, KV>::cvt(c[j], sc);
          constexpr int OUT_VECS = 16 * (int)sizeof(T) / 16;
#pragma unroll
          for (int j = 0; j < OUT_VECS; ++j)
            reinterpret_cast<uint4*>(d + 16 * i)[j] = reinterpret_cast<const uint4*>(o)[j];
        }
      }
    } else {
      for (int i = lane; i < entry_size; i += 32) d[i] = GatherCvt<ST, KV>::cvt(s[i], sc);
    }
  }
}

// cp_gather_cache: raw copy of whole entries, any element width (the reference dispatches on
// 8/16/32-bit words, cache_kernels.cu:911-925); seq_starts counts TOKENS and need not be a
// multiple of block_size (:816-823).  All sizes and strides in BYTES.
template <typename V>
__global__ void __launch_bounds__(kCopyThreads)
cp_gather_cache_kernel(const char* __restrict__ src_cache, char* __restrict__ dst,
                       const int32_t* __restrict__ block_table,
                       const int32_t* __restrict__ cu_seq_lens, int batch_size, int block_size,
                       int entry_bytes, int64_t block_table_stride, int64_t cache_block_bytes,
                       int64_t cache_entry_bytes, int64_t dst_entry_bytes,
                       const int32_t* __restrict__ seq_starts) {
  const int lane = threadIdx.x & 31;
  const int warps = gridDim.x * (kCopyThreads / 32);
  const int total = cu_seq_lens[batch_size];
  const int n = entry_bytes / (int)sizeof(V);
  for (int tok = blockIdx.x * (kCopyThreads / 32) + (threadIdx.x >> 5); tok < total; tok += warps) {
    const int b = seq_of_token(cu_seq_lens, batch_size, tok);
    int pos = tok - cu_seq_lens[b];
    if (seq_starts != nullptr) pos += seq_starts[b];
    const int page = pos / block_size, slot = pos - page * block_size;
    const int64_t block_id = block_table[(int64_t)b * block_table_stride + page];
    const V* s = reinterpret_cast<const V*>(src_cache + block_id * cache_block_bytes +
                                            (int64_t)slot * cache_entry_bytes);
    V* d = reinterpret_cast<V*>(dst + (int64_t)tok * dst_entry_bytes);
    for (int i = lane; i < n; i += 32) d[i] = s[i];
  }
}

static dim3 gather_grid(int64_t tot_tokens_hint, int num_sms) {
  // enough warps to cover the tokens once, capped at a full machine of resident CTAs
  const int wpb = kCopyThreads / 32;
  int64_t blocks = (tot_tokens_hint + wpb - 1) / wpb;
  const int64_t cap = (int64_t)num_sms * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return dim3((unsigned)blocks);
}

template <int ST, int KV>
int launch_gather_t(const void* src_cache, void* dst, const int32_t* block_table,
                    const int32_t* cu_seq_lens, int batch_size, int block_size, int entry_size,
                    int64_t block_table_stride, int64_t cache_block_stride,
                    int64_t cache_entry_stride, int64_t dst_entry_stride, const float* scale,
                    const int32_t* seq_starts, int64_t tot_tokens_hint, int num_sms,
                    cudaStream_t stream) {
  using S = typename GatherCvt<ST, KV>::src_t;
  using T = typename Scalar<ST>::type;
  const int64_t ses = sizeof(S), des = sizeof(T);
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
  const bool vec = al16(src_cache) && al16(dst) && (cache_block_stride * ses) % 16 == 0 &&
                   (cache_entry_stride * ses) % 16 == 0 && (dst_entry_stride * des) % 16 == 0 &&
                   (KV == B200PA_KV_AUTO ? (entry_size * des) % 16 == 0 : entry_size % 16 == 0);
  const dim3 grid = gather_grid(tot_tokens_hint, num_sms), block(kCopyThreads);
#define GATHER_CALL(VEC)                                                                        \
  gather_cache_kernel<ST, KV, VEC><<<grid, block, 0, stream>>>(                             \
      (const S*)src_cache, (T*)dst, block_table, cu_seq_lens, batch_size, block_size, entry_size, \
      block_table_stride, cache_block_stride, cache_entry_stride, dst_entry_stride, scale,      \
      seq_starts)
  if (vec) GATHER_CALL(true); else GATHER_CALL(false);
#undef GATHER_CALL
  count_launch();
  return check_launch("gather_and_maybe_dequant_cache");
}

}  // namespace

int launch_copy_blocks(void* const* key_ptrs, void* const* value_ptrs, int num_layers,
                       const int64_t* block_mapping, int num_pairs, int64_t block_bytes,
                       cudaStream_t stream) {
  bool al16 = block_bytes % 16 == 0;
  for (int l = 0; l < num_layers && al16; ++l)
    al16 = (reinterpret_cast<uintptr_t>(key_ptrs[l]) % 16 == 0) &&
           (reinterpret_cast<uintptr_t>(value_ptrs[l]) % 16 == 0);
  for (int l0 = 0; l0 < num_layers; l0 += kPtrChunk) {
    const int nl = num_layers - l0 < kPtrChunk ? num_layers - l0 : kPtrChunk;
    LayerPtrs ptrs;
    for (int l = 0; l < nl; ++l) {
      ptrs.key[l] = key_ptrs[l0 + l];
      ptrs.value[l] = value_ptrs[l0 + l];
    }
    // grid.y is limited to 65535 pairs per launch
    for (int p0 = 0; p0 < num_pairs; p0 += 65535) {
      const int np = num_pairs - p0 < 65535 ? num_pairs - p0 : 65535;
      const dim3 grid(nl, np), block(kCopyThreads);
      if (al16)
        copy_blocks_kernel<uint4><<<grid, block, 0, stream>>>(ptrs, block_mapping + 2 * (int64_t)p0, block_bytes);
      else
        copy_blocks_kernel<uint8_t><<<grid, block, 0, stream>>>(ptrs, block_mapping + 2 * (int64_t)p0, block_bytes);
      count_launch();
    }
  }
  return check_launch("copy_blocks");
}

int launch_gather_cache(const void* src_cache, void* dst, const int32_t* block_table,
                        const int32_t* cu_seq_lens, int batch_size, int block_size, int entry_size,
                        int64_t block_table_stride, int64_t cache_block_stride,
                        int64_t cache_entry_stride, int64_t dst_entry_stride, int scalar_type,
                        int kv_dtype, const float* scale, const int32_t* seq_starts,
                        int64_t tot_tokens_hint, int num_sms, cudaStream_t stream) {
#define GATHER_ARGS                                                                               \
  src_cache, dst, block_table, cu_seq_lens, batch_size, block_size, entry_size, block_table_stride, \
      cache_block_stride, cache_entry_stride, dst_entry_stride, scale, seq_starts, tot_tokens_hint, \
      num_sms, stream
#define GATHER_KV(ST)                                                                              \
  if (kv_dtype == B200PA_KV_AUTO) return launch_gather_t<ST, B200PA_KV_AUTO>(GATHER_ARGS);         \
  if (kv_dtype == B200PA_KV_FP8_E4M3) return launch_gather_t<ST, B200PA_KV_FP8_E4M3>(GATHER_ARGS); \
  return launch_gather_t<ST, B200PA_KV_FP8_E5M2>(GATHER_ARGS)
  if (scalar_type == B200PA_BF16) { GATHER_KV(B200PA_BF16); }
  if (scalar_type == B200PA_F16) { GATHER_KV(B200PA_F16); }
  GATHER_KV(B200PA_F32);
#undef GATHER_KV
#undef GATHER_ARGS
}

// element_bytes in {1, 2, 4}; sizes / strides in ELEMENTS as the reference passes them
int launch_cp_gather_cache(const void* src_cache, void* dst, const int32_t* block_table,
                           const int32_t* cu_seq_lens, int batch_size, int block_size,
                           int entry_size, int64_t block_table_stride, int64_t cache_block_stride,
                           int64_t cache_entry_stride, int64_t dst_entry_stride, int element_bytes,
                           const int32_t* seq_starts, int64_t tot_tokens_hint, int num_sms,
                           cudaStream_t stream) {
  const int64_t eb = element_bytes;
  const int entry_bytes = (int)(entry_size * eb);
  const int64_t cbb = cache_block_stride * eb, ceb = cache_entry_stride * eb, deb = dst_entry_stride * eb;
  auto al = [](const void* p, uintptr_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; };
  const dim3 grid = gather_grid(tot_tokens_hint, num_sms), block(kCopyThreads);
#define CP_CALL(V)                                                                             \
  cp_gather_cache_kernel<V><<<grid, block, 0, stream>>>(                                       \
      (const char*)src_cache, (char*)dst, block_table, cu_seq_lens, batch_size, block_size,    \
      entry_bytes, block_table_stride, cbb, ceb, deb, seq_starts)
  auto fits = [&](int64_t a) {
    return al(src_cache, a) && al(dst, a) && cbb % a == 0 && ceb % a == 0 && deb % a == 0 &&
           entry_bytes % a == 0;
  };
  if (fits(16)) CP_CALL(uint4);
  else if (fits(4)) CP_CALL(uint32_t);
  else if (fits(2)) CP_CALL(uint16_t);
  else CP_CALL(uint8_t);
#undef CP_CALL
  count_launch();
  return check_launch("cp_gather_cache");
}

}  // namespace b200pa
