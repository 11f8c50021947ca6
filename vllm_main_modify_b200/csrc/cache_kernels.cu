// KV-cache write / quantise kernels for sm_100a.
//
// Semantics follow the reference ops (paths relative to the reference root):
//   reshape_and_cache        csrc/cache_kernels.cu:212-264, 416-442
//   reshape_and_cache_flash  csrc/cache_kernels.cu:280-353, 459-497
//   convert_fp8              csrc/cache_kernels.cu:552-623
// The quantise primitive is fp8(float(x) / scale) with IEEE division, RNE,
// saturate-to-finite (csrc/quantization/fp8/nvidia/quant_utils.cuh:469-502), so
// the produced bytes are bit-identical to the reference's.
//
// These are HBM/latency-bound scatter kernels (a few KiB per token): the design
// goal is coalesced 16-byte loads of the token row, 16-byte stores wherever the
// destination layout has 16 contiguous bytes, and one launch for K and V.
#include "common.cuh"

namespace b200pa {

namespace {

template <int ST> struct ElemBytes { static constexpr int v = (ST == B200PA_F32) ? 4 : 2; };

// Load `N` consecutive scalar_t as floats (generic, any alignment).
template <int ST, int N>
__device__ __forceinline__ void load_floats(const typename Scalar<ST>::type* p, float* f) {
#pragma unroll
  for (int i = 0; i < N; ++i) f[i] = Scalar<ST>::to_float(p[i]);
}

// ---------------------------------------------------------------------------
// reshape_and_cache: x-packed K [NB, H, D/x, BS, x], V [NB, H, D, BS].
// One CTA per token.  K is written x elements (= 16 bytes) per thread; V is an
// inherent stride-BS scatter (one element per (head, d) row).
// ---------------------------------------------------------------------------
template <int ST, int KV, bool VEC>
__global__ void reshape_and_cache_kernel(
    const typename Scalar<ST>::type* __restrict__ key,
    const typename Scalar<ST>::type* __restrict__ value, void* __restrict__ key_cache,
    void* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping,
    int64_t key_stride, int64_t value_stride, int num_heads, int head_size,
    int block_size, int x, const float* __restrict__ k_scale,
    const float* __restrict__ v_scale) {
  using T = typename Scalar<ST>::type;
  // programmatic dependent launch: a decode kernel queued behind this one (launched with
  // programmatic stream serialisation) may be scheduled now; it blocks in griddepcontrol.wait
  // until this grid has completed and its stores are visible.  A no-op for any other successor.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;  // padding token
  const int64_t block_idx = slot / block_size;
  const int64_t block_off = slot % block_size;
  const int n = num_heads * head_size;
  const T* ksrc = key + token * key_stride;
  const T* vsrc = value + token * value_stride;
  const int64_t blk_elems = (int64_t)num_heads * head_size * block_size;

  float ks = 1.f, vs = 1.f;
  if constexpr (KV != B200PA_KV_AUTO) { ks = *k_scale; vs = *v_scale; }

  // ---- K: chunks of x elements -> 16 contiguous bytes in the cache
  const int chunks_per_head = head_size / x;
  const int n_chunks = num_heads * chunks_per_head;
  for (int c = threadIdx.x; c < n_chunks; c += blockDim.x) {
    const int head = c / chunks_per_head;
    const int xi = c - head * chunks_per_head;
    const T* src = ksrc + head * head_size + xi * x;
    const int64_t dst = block_idx * blk_elems + (int64_t)head * head_size * block_size +
                        (int64_t)xi * block_size * x + block_off * x;
    if constexpr (KV == B200PA_KV_AUTO) {
      T* kc = reinterpret_cast<T*>(key_cache) + dst;
      if constexpr (VEC) {
        *reinterpret_cast<uint4*>(kc) = *reinterpret_cast<const uint4*>(src);
      } else {
        for (int i = 0; i < x; ++i) kc[i] = src[i];
      }
    } else {
      uint8_t* kc = reinterpret_cast<uint8_t*>(key_cache) + dst;
      if constexpr (VEC && ST != B200PA_F32) {  // x == 16: 32 bytes in, 16 bytes out
        uint4 raw[2];
        raw[0] = reinterpret_cast<const uint4*>(src)[0];
        raw[1] = reinterpret_cast<const uint4*>(src)[1];
        const T* e = reinterpret_cast<const T*>(raw);
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t lo = fp8_quant2<KV>(Scalar<ST>::to_float(e[4 * i + 0]),
                                             Scalar<ST>::to_float(e[4 * i + 1]), ks);
          const uint32_t hi = fp8_quant2<KV>(Scalar<ST>::to_float(e[4 * i + 2]),
                                             Scalar<ST>::to_float(e[4 * i + 3]), ks);
          w[i] = lo | (hi << 16);
        }
        *reinterpret_cast<uint4*>(kc) = make_uint4(w[0], w[1], w[2], w[3]);
      } else {
        for (int i = 0; i < x; ++i) kc[i] = fp8_quant<KV>(Scalar<ST>::to_float(src[i]), ks);
      }
    }
  }

  // ---- V: one element per (head, d) row
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int64_t dst = block_idx * blk_elems + (int64_t)i * block_size + block_off;
    if constexpr (KV == B200PA_KV_AUTO) {
      reinterpret_cast<T*>(value_cache)[dst] = vsrc[i];
    } else {
      reinterpret_cast<uint8_t*>(value_cache)[dst] =
          fp8_quant<KV>(Scalar<ST>::to_float(vsrc[i]), vs);
    }
  }
}

// ---------------------------------------------------------------------------
// reshape_and_cache for prefill-sized batches: one CTA per (tile of kRunTokens consecutive
// tokens, kv head).  Same bytes as the kernel above.  When the tile's tokens are consecutive
// positions of one sequence (slot offsets advance by one, a new physical block at every multiple
// of block_size) -- the shape of every prefill chunk -- the V rows of the tile are staged in
// shared memory TRANSPOSED, laid out exactly as the cache rows they go to ([d][block j][token],
// the first token at its slot offset), and copied out with 16-byte stores: the cache's
// [D][block_size] plane of a full block is one contiguous run.  The per-token kernel writes the
// same plane one 1-2 byte element at a stride of block_size (cache_kernels.cu:254-262 of the
// reference does too).  Tiles that are not such a run (decode tokens, a sequence boundary, a
// padding slot) take the per-element path inside this kernel.
// ---------------------------------------------------------------------------
constexpr int kRunTokens = 64;
constexpr int kRunThreads = 256;
constexpr int kRunMinTokens = 1024;    // below this the per-token kernel has more CTAs in flight

template <int ST, int KV>
__global__ void __launch_bounds__(kRunThreads)
reshape_and_cache_run_kernel(
    const typename Scalar<ST>::type* __restrict__ key,
    const typename Scalar<ST>::type* __restrict__ value, void* __restrict__ key_cache,
    void* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping, int num_tokens,
    int64_t key_stride, int64_t value_stride, int num_heads, int head_size, int block_size,
    const float* __restrict__ k_scale, const float* __restrict__ v_scale) {
  using T = typename Scalar<ST>::type;
  static_assert(ST != B200PA_F32, "16-bit sources");
  constexpr int CES = KV == B200PA_KV_AUTO ? 2 : 1;     // cache element bytes
  constexpr int X = 16 / CES;                            // elements per 16-byte K chunk
  extern __shared__ __align__(16) uint8_t vs_smem[];     // [head_size][pitch] V rows as in the cache
  __shared__ int blk_s[kRunTokens];    // physical block of the token, -1: padding slot
  __shared__ int off_s[kRunTokens];    // slot offset in its block
  __shared__ int run_s;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int t0 = blockIdx.x * kRunTokens;
  const int head = blockIdx.y;
  const int nt = min(kRunTokens, num_tokens - t0);
  if (threadIdx.x < kRunTokens) {
    const int64_t sl = threadIdx.x < nt ? slot_mapping[t0 + threadIdx.x] : -1;
    blk_s[threadIdx.x] = sl < 0 ? -1 : (int)(sl / block_size);
    off_s[threadIdx.x] = sl < 0 ? 0 : (int)(sl % block_size);
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    // a run: every slot valid; offsets advance by one; inside a block the block stays the same
    bool ok = true;
    for (int i = threadIdx.x; i < nt; i += 32) {
      ok = ok && blk_s[i] >= 0;
      if (i > 0) {
        const int want = off_s[i - 1] + 1 == block_size ? 0 : off_s[i - 1] + 1;
        ok = ok && off_s[i] == want && (want == 0 || blk_s[i] == blk_s[i - 1]);
      }
    }
    ok = __all_sync(0xffffffffu, ok);
    if (threadIdx.x == 0) run_s = ok ? 1 : 0;
  }
  __syncthreads();
  const bool run = run_s != 0;
  float ks = 1.f, vs = 1.f;
  if constexpr (KV != B200PA_KV_AUTO) { ks = *k_scale; vs = *v_scale; }
  const int64_t blk_bytes = (int64_t)num_heads * head_size * block_size * CES;
  const int64_t head_bytes = (int64_t)head * head_size * block_size * CES;

  // ---- K: one 16-byte cache chunk per (token, chunk), lanes along the chunks of a token row
  const int cph = head_size / X;
#pragma unroll 2
  for (int idx = threadIdx.x; idx < nt * cph; idx += kRunThreads) {
    const int i = idx / cph, xi = idx - i * cph;
    const int blk = blk_s[i];
    if (blk < 0) continue;
    const T* src = key + (int64_t)(t0 + i) * key_stride + head * head_size + xi * X;
    uint8_t* dst = reinterpret_cast<uint8_t*>(key_cache) + blk * blk_bytes + head_bytes +
                   ((int64_t)xi * block_size + off_s[i]) * 16;
    if constexpr (KV == B200PA_KV_AUTO) {
      *reinterpret_cast<uint4*>(dst) = __ldg(reinterpret_cast<const uint4*>(src));
    } else {
      uint4 raw[2];
      raw[0] = __ldg(reinterpret_cast<const uint4*>(src));
      raw[1] = __ldg(reinterpret_cast<const uint4*>(src) + 1);
      const T* e = reinterpret_cast<const T*>(raw);
      uint32_t w[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t lo = fp8_quant2<KV>(Scalar<ST>::to_float(e[4 * k + 0]),
                                           Scalar<ST>::to_float(e[4 * k + 1]), ks);
        const uint32_t hi = fp8_quant2<KV>(Scalar<ST>::to_float(e[4 * k + 2]),
                                           Scalar<ST>::to_float(e[4 * k + 3]), ks);
        w[k] = lo | (hi << 16);
      }
      *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }

  // ---- V
  const int d8n = head_size / 8;                       // 16-byte groups of a source row
  if (!run) {
    // no run of positions: the per-token kernel's pattern -- lanes along the head elements
    // (coalesced loads; a warp's stores fall into 8 cache lines), four passes in flight
    const int total = nt * head_size;
    for (int base = 0; base < total; base += 4 * kRunThreads) {
      T v[4];
      int64_t dsto[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int idx = base + u * kRunThreads + threadIdx.x;
        dsto[u] = -1;
        if (idx < total) {
          const int i = idx / head_size, d = idx - i * head_size;
          const int blk = blk_s[i];
          if (blk >= 0) {
            v[u] = value[(int64_t)(t0 + i) * value_stride + head * head_size + d];
            dsto[u] = blk * blk_bytes + head_bytes + ((int64_t)d * block_size + off_s[i]) * CES;
          }
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (dsto[u] >= 0) {
          uint8_t* dst = reinterpret_cast<uint8_t*>(value_cache) + dsto[u];
          if constexpr (KV == B200PA_KV_AUTO) *reinterpret_cast<T*>(dst) = v[u];
          else *dst = fp8_quant<KV>(Scalar<ST>::to_float(v[u]), vs);
        }
      }
    }
    return;
  }
  const int a = off_s[0];                              // slot offset of the tile's first token
  const int pitch = (kRunTokens / block_size + 1) * block_size * CES;   // staged row, bytes
  // stage: lanes along the tokens (conflict-free 1-2 byte shared stores), 8 head elements each
  for (int idx = threadIdx.x; idx < d8n * kRunTokens; idx += kRunThreads) {
    const int i = idx % kRunTokens, d8 = idx / kRunTokens;
    if (i >= nt) continue;
    const uint4 raw = __ldg(reinterpret_cast<const uint4*>(
        value + (int64_t)(t0 + i) * value_stride + head * head_size + d8 * 8));
    const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      uint8_t* dst = vs_smem + (size_t)(d8 * 8 + k) * pitch + (a + i) * CES;
      if constexpr (KV == B200PA_KV_AUTO) *reinterpret_cast<T*>(dst) = e[k];
      else *dst = fp8_quant<KV>(Scalar<ST>::to_float(e[k]), vs);
    }
  }
  __syncthreads();
  // copy out: (block j, row d, 16-byte group g) -- for a full block the rows of one head are one
  // contiguous run of head_size * block_size elements
  const int gpb = block_size * CES / 16 > 0 ? block_size * CES / 16 : 1;   // groups per block row
  const int gbytes = block_size * CES < 16 ? block_size * CES : 16;          // (block_size 8, fp8: 8)
  const int nblk = (a + nt + block_size - 1) / block_size;
  const int lo_b = a * CES, hi_b = (a + nt) * CES;     // valid byte range of a staged row
  const int per_blk = head_size * gpb;
  for (int idx = threadIdx.x; idx < nblk * per_blk; idx += kRunThreads) {
    const int j = idx / per_blk;
    const int rem = idx - j * per_blk;
    const int d = rem / gpb, g = rem - d * gpb;
    const int tok = max(0, j * block_size - a);        // first token of the tile in block j
    const int b0 = j * block_size * CES + g * gbytes;  // byte offset in the staged row
    const uint8_t* src = vs_smem + (size_t)d * pitch + b0;
    uint8_t* dst = reinterpret_cast<uint8_t*>(value_cache) + blk_s[tok] * blk_bytes + head_bytes +
                   (int64_t)d * block_size * CES + g * gbytes;
    if (gbytes == 16 && b0 >= lo_b && b0 + 16 <= hi_b) {
      *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
    } else {
      for (int k = 0; k < gbytes; k += CES) {
        if (b0 + k >= lo_b && b0 + k < hi_b) {
          if constexpr (CES == 2) *reinterpret_cast<uint16_t*>(dst + k) = *reinterpret_cast<const uint16_t*>(src + k);
          else dst[k] = src[k];
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// reshape_and_cache_flash: NHD / HND, addressed by strides.  One CTA per token;
// each thread moves 8 consecutive elements of one head (16-byte load).
// ---------------------------------------------------------------------------
template <int ST, int KV, bool VEC>
__global__ void reshape_and_cache_flash_kernel(
    const typename Scalar<ST>::type* __restrict__ key,
    const typename Scalar<ST>::type* __restrict__ value, void* __restrict__ key_cache,
    void* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping,
    int64_t key_stride, int64_t value_stride, int64_t block_stride,
    int64_t page_stride, int64_t head_stride, int num_heads, int head_size,
    int block_size, const float* __restrict__ k_scale,
    const float* __restrict__ v_scale) {
  using T = typename Scalar<ST>::type;
  using C = typename std::conditional<KV == B200PA_KV_AUTO, T, uint8_t>::type;
  // programmatic dependent launch: a decode kernel queued behind this one (launched with
  // programmatic stream serialisation) may be scheduled now; it blocks in griddepcontrol.wait
  // until this grid has completed and its stores are visible.  A no-op for any other successor.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;
  const int64_t block_idx = slot / block_size;
  const int64_t block_off = slot % block_size;
  const T* ksrc = key + token * key_stride;
  const T* vsrc = value + token * value_stride;
  C* kdst = reinterpret_cast<C*>(key_cache) + block_idx * block_stride + block_off * page_stride;
  C* vdst = reinterpret_cast<C*>(value_cache) + block_idx * block_stride + block_off * page_stride;
  float ks = 1.f, vs = 1.f;
  if constexpr (KV != B200PA_KV_AUTO) { ks = *k_scale; vs = *v_scale; }

  constexpr int V = VEC ? (16 / ElemBytes<ST>::v) : 1;  // elements per thread step
  const int per_head = head_size / V;
  const int n_vec = num_heads * per_head;
  for (int i = threadIdx.x; i < 2 * n_vec; i += blockDim.x) {
    const bool is_v = i >= n_vec;
    const int j = is_v ? i - n_vec : i;
    const int head = j / per_head;
    const int d = (j - head * per_head) * V;
    const T* src = (is_v ? vsrc : ksrc) + head * head_size + d;
    C* dst = (is_v ? vdst : kdst) + (int64_t)head * head_stride + d;
    const float sc = is_v ? vs : ks;
    if constexpr (VEC) {
      const uint4 raw = *reinterpret_cast<const uint4*>(src);
      if constexpr (KV == B200PA_KV_AUTO) {
        *reinterpret_cast<uint4*>(dst) = raw;
      } else {
        const T* e = reinterpret_cast<const T*>(&raw);
        if constexpr (V == 8) {
          uint32_t w[2];
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            const uint32_t lo = fp8_quant2<KV>(Scalar<ST>::to_float(e[4 * k + 0]),
                                               Scalar<ST>::to_float(e[4 * k + 1]), sc);
            const uint32_t hi = fp8_quant2<KV>(Scalar<ST>::to_float(e[4 * k + 2]),
                                               Scalar<ST>::to_float(e[4 * k + 3]), sc);
            w[k] = lo | (hi << 16);
          }
          *reinterpret_cast<uint2*>(dst) = make_uint2(w[0], w[1]);
        } else {  // fp32 source: 4 elements
          const uint32_t lo = fp8_quant2<KV>(Scalar<ST>::to_float(e[0]), Scalar<ST>::to_float(e[1]), sc);
          const uint32_t hi = fp8_quant2<KV>(Scalar<ST>::to_float(e[2]), Scalar<ST>::to_float(e[3]), sc);
          *reinterpret_cast<uint32_t*>(dst) = lo | (hi << 16);
        }
      }
    } else {
      if constexpr (KV == B200PA_KV_AUTO) {
        dst[0] = src[0];
      } else {
        dst[0] = fp8_quant<KV>(Scalar<ST>::to_float(src[0]), sc);
      }
    }
  }
}

// ---------------------------------------------------------------------------
// convert_fp8: whole-tensor fp8 <-> scalar_t with a host scale (test utility).
// ---------------------------------------------------------------------------
template <int ST, int KV, bool TO_FP8>
__global__ void convert_fp8_kernel(void* __restrict__ dst, const void* __restrict__ src,
                                   int64_t n, float scale) {
  using T = typename Scalar<ST>::type;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if constexpr (TO_FP8) {
      reinterpret_cast<uint8_t*>(dst)[i] =
          fp8_quant<KV>(Scalar<ST>::to_float(reinterpret_cast<const T*>(src)[i]), scale);
    } else {
      const float v = fp8_to_float<KV>(reinterpret_cast<const uint8_t*>(src)[i]) * scale;
      reinterpret_cast<T*>(dst)[i] = Scalar<ST>::from_float(v);
    }
  }
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

#define B200PA_DISPATCH_ST_KV(ST_VAL, KV_VAL, ...)                                 \
  do {                                                                               \
    if ((ST_VAL) == B200PA_BF16) {                                                   \
      constexpr int ST = B200PA_BF16;                                                \
      if ((KV_VAL) == B200PA_KV_AUTO) { constexpr int KV = B200PA_KV_AUTO; __VA_ARGS__; }   \
      else if ((KV_VAL) == B200PA_KV_FP8_E4M3) { constexpr int KV = B200PA_KV_FP8_E4M3; __VA_ARGS__; } \
      else { constexpr int KV = B200PA_KV_FP8_E5M2; __VA_ARGS__; }                          \
    } else if ((ST_VAL) == B200PA_F16) {                                             \
      constexpr int ST = B200PA_F16;                                                 \
      if ((KV_VAL) == B200PA_KV_AUTO) { constexpr int KV = B200PA_KV_AUTO; __VA_ARGS__; }   \
      else if ((KV_VAL) == B200PA_KV_FP8_E4M3) { constexpr int KV = B200PA_KV_FP8_E4M3; __VA_ARGS__; } \
      else { constexpr int KV = B200PA_KV_FP8_E5M2; __VA_ARGS__; }                          \
    } else {                                                                         \
      constexpr int ST = B200PA_F32;                                                 \
      if ((KV_VAL) == B200PA_KV_AUTO) { constexpr int KV = B200PA_KV_AUTO; __VA_ARGS__; }   \
      else if ((KV_VAL) == B200PA_KV_FP8_E4M3) { constexpr int KV = B200PA_KV_FP8_E4M3; __VA_ARGS__; } \
      else { constexpr int KV = B200PA_KV_FP8_E5M2; __VA_ARGS__; }                          \
    }                                                                                \
  } while (0)

int launch_reshape_and_cache(const void* key, const void* value, void* key_cache,
                             void* value_cache, const int64_t* slot_mapping,
                             int num_tokens, int64_t key_stride, int64_t value_stride,
                             int num_heads, int head_size, int block_size, int x,
                             int scalar_type, int kv_dtype, const float* k_scale,
                             const float* v_scale, cudaStream_t stream) {
  if (num_tokens == 0) return B200PA_OK;
  const int es = scalar_type == B200PA_F32 ? 4 : 2;
  const int cache_es = kv_dtype == B200PA_KV_AUTO ? es : 1;
  if (x <= 0 || head_size % x != 0)
    return set_error(B200PA_ERR_INVALID, "reshape_and_cache: head_size %d not divisible by x %d", head_size, x);
  // 16-byte vector path: x elements are exactly 16 cache bytes and the source
  // chunk is 16-byte aligned for every token / head / chunk.
  const bool vec = (x * cache_es == 16) && aligned16(key) && aligned16(key_cache) &&
                   ((key_stride * es) % 16 == 0) && ((head_size * es) % 16 == 0) &&
                   ((x * es) % 16 == 0) && (scalar_type != B200PA_F32 || kv_dtype == B200PA_KV_AUTO);
  // prefill-sized batches: token tiles x kv heads, V staged and stored by rows (see the kernel)
  const bool run_ok = vec && num_tokens >= kRunMinTokens && scalar_type != B200PA_F32 &&
                      aligned16(value) && aligned16(value_cache) &&
                      ((value_stride * es) % 16 == 0) && head_size % 8 == 0 &&
                      (block_size == 8 || block_size == 16 || block_size == 32) &&
                      num_heads <= 65535 &&
                      head_size * (kRunTokens / block_size + 1) * block_size * cache_es <= 47104;
  if (run_ok) {
    const int pitch = (kRunTokens / block_size + 1) * block_size * cache_es;
    const int smem = head_size * pitch;
    dim3 grid((num_tokens + kRunTokens - 1) / kRunTokens, num_heads);
#define RUN_CASE(ST_, KV_)                                                                     \
    reshape_and_cache_run_kernel<ST_, KV_><<<grid, kRunThreads, smem, stream>>>(               \
        (const typename Scalar<ST_>::type*)key, (const typename Scalar<ST_>::type*)value,      \
        key_cache, value_cache, slot_mapping, num_tokens, key_stride, value_stride, num_heads, \
        head_size, block_size, k_scale, v_scale)
    if (scalar_type == B200PA_BF16) {
      if (kv_dtype == B200PA_KV_AUTO) RUN_CASE(B200PA_BF16, B200PA_KV_AUTO);
      else if (kv_dtype == B200PA_KV_FP8_E4M3) RUN_CASE(B200PA_BF16, B200PA_KV_FP8_E4M3);
      else RUN_CASE(B200PA_BF16, B200PA_KV_FP8_E5M2);
    } else {
      if (kv_dtype == B200PA_KV_AUTO) RUN_CASE(B200PA_F16, B200PA_KV_AUTO);
      else if (kv_dtype == B200PA_KV_FP8_E4M3) RUN_CASE(B200PA_F16, B200PA_KV_FP8_E4M3);
      else RUN_CASE(B200PA_F16, B200PA_KV_FP8_E5M2);
    }
#undef RUN_CASE
    count_launch();
    return check_launch("reshape_and_cache");
  }
  const int threads = min(max(num_heads * head_size / 4, 64), 512);
  dim3 grid(num_tokens), block((threads + 31) / 32 * 32);
  B200PA_DISPATCH_ST_KV(scalar_type, kv_dtype, {
    using T = typename Scalar<ST>::type;
    if (vec) {
      reshape_and_cache_kernel<ST, KV, true><<<grid, block, 0, stream>>>(
          (const T*)key, (const T*)value, key_cache, value_cache, slot_mapping, key_stride,
          value_stride, num_heads, head_size, block_size, x, k_scale, v_scale);
    } else {
      reshape_and_cache_kernel<ST, KV, false><<<grid, block, 0, stream>>>(
          (const T*)key, (const T*)value, key_cache, value_cache, slot_mapping, key_stride,
          value_stride, num_heads, head_size, block_size, x, k_scale, v_scale);
    }
  });
  count_launch();
  return check_launch("reshape_and_cache");
}

int launch_reshape_and_cache_flash(const void* key, const void* value, void* key_cache,
                                   void* value_cache, const int64_t* slot_mapping,
                                   int num_tokens, int64_t key_stride, int64_t value_stride,
                                   int64_t block_stride, int64_t page_stride,
                                   int64_t head_stride, int num_heads, int head_size,
                                   int block_size, int scalar_type, int kv_dtype,
                                   const float* k_scale, const float* v_scale,
                                   cudaStream_t stream) {
  if (num_tokens == 0) return B200PA_OK;
  const int es = scalar_type == B200PA_F32 ? 4 : 2;
  const int ces = kv_dtype == B200PA_KV_AUTO ? es : 1;
  const int V = 16 / es;
  const bool vec = (head_size % V == 0) && aligned16(key) && aligned16(value) &&
                   ((key_stride * es) % 16 == 0) && ((value_stride * es) % 16 == 0) &&
                   ((head_size * es) % 16 == 0) &&
                   ((reinterpret_cast<uintptr_t>(key_cache) % (V * ces)) == 0) &&
                   ((reinterpret_cast<uintptr_t>(value_cache) % (V * ces)) == 0) &&
                   ((block_stride * ces) % (V * ces) == 0) && ((page_stride * ces) % (V * ces) == 0) &&
                   ((head_stride * ces) % (V * ces) == 0);
  const int work = 2 * num_heads * head_size / (vec ? V : 1);
  const int threads = min(max((work + 31) / 32 * 32, 32), 512);
  dim3 grid(num_tokens), block(threads);
  B200PA_DISPATCH_ST_KV(scalar_type, kv_dtype, {
    using T = typename Scalar<ST>::type;
    if (vec) {
      reshape_and_cache_flash_kernel<ST, KV, true><<<grid, block, 0, stream>>>(
          (const T*)key, (const T*)value, key_cache, value_cache, slot_mapping, key_stride,
          value_stride, block_stride, page_stride, head_stride, num_heads, head_size,
          block_size, k_scale, v_scale);
    } else {
      reshape_and_cache_flash_kernel<ST, KV, false><<<grid, block, 0, stream>>>(
          (const T*)key, (const T*)value, key_cache, value_cache, slot_mapping, key_stride,
          value_stride, block_stride, page_stride, head_stride, num_heads, head_size,
          block_size, k_scale, v_scale);
    }
  });
  count_launch();
  return check_launch("reshape_and_cache_flash");
}

int launch_convert_fp8(void* dst, const void* src, int64_t n, float scale, int scalar_type,
                       int kv_dtype, int to_fp8, cudaStream_t stream) {
  if (n == 0) return B200PA_OK;
  if (kv_dtype != B200PA_KV_FP8_E4M3 && kv_dtype != B200PA_KV_FP8_E5M2)
    return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported data type: convert_fp8 needs fp8 / fp8_e4m3 / fp8_e5m2");
  const int threads = 256;
  const int64_t want = (n + threads - 1) / threads;
  const int blocks = (int)(want < 148 * 16 ? want : 148 * 16);
#define CONV(ST_, KV_)                                                                      \
  do {                                                                                      \
    if (to_fp8) convert_fp8_kernel<ST_, KV_, true><<<blocks, threads, 0, stream>>>(dst, src, n, scale); \
    else convert_fp8_kernel<ST_, KV_, false><<<blocks, threads, 0, stream>>>(dst, src, n, scale);       \
  } while (0)
  if (kv_dtype == B200PA_KV_FP8_E4M3) {
    if (scalar_type == B200PA_BF16) CONV(B200PA_BF16, B200PA_KV_FP8_E4M3);
    else if (scalar_type == B200PA_F16) CONV(B200PA_F16, B200PA_KV_FP8_E4M3);
    else CONV(B200PA_F32, B200PA_KV_FP8_E4M3);
  } else {
    if (scalar_type == B200PA_BF16) CONV(B200PA_BF16, B200PA_KV_FP8_E5M2);
    else if (scalar_type == B200PA_F16) CONV(B200PA_F16, B200PA_KV_FP8_E5M2);
    else CONV(B200PA_F32, B200PA_KV_FP8_E5M2);
  }
#undef CONV
  count_launch();
  return check_launch("convert_fp8");
}

}  // namespace b200pa
