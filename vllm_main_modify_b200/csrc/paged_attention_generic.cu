// Generic paged-decode attention kernel (SIMT) + the split-K reduce kernel.
//
// Covers the WHOLE domain of the reference ops (paged_attention_v1.cu:90-158 of
// the reference): head sizes {32,64,80,96,112,120,128,192,256}, block sizes
// {8,16,32}, query dtype {fp32,fp16,bf16}, kv {auto, fp8_e4m3, fp8_e5m2}, ALiBi
// and block-sparse.  The shapes of the BASELINE configs are served by the
// persistent stream-K tensor-core kernel in paged_attention_stream.cu; this kernel is the
// always-available companion (and the in-GPU cross-check in the tests).
//
// Differences from the reference kernel (attention_kernels.cuh:85-491) are
// structural, not semantic: online softmax per warp (no context-length shared
// memory, so no max_seq_len limit), fp32 un-normalised probabilities for P.V,
// coalesced 16-byte loads of V rows.  fp8 values are dequantised bit-faithfully
// (fp8 -> half -> float * scale -> scalar_t), quant_utils.cuh:306-428.
#include "common.cuh"

namespace b200pa {

namespace {

constexpr int kGenericThreads = 128;
constexpr int kGenericWarps = kGenericThreads / kWarp;
constexpr int kMaxHead = 256;
constexpr int kMaxRows = kMaxHead / kWarp;  // V rows (head elements) per lane

template <int ST, int KV> struct CacheElem {
  using type = typename std::conditional<KV == B200PA_KV_AUTO, typename Scalar<ST>::type, uint8_t>::type;
};

template <int ST, int KV>
__device__ __forceinline__ float cache_to_float(typename CacheElem<ST, KV>::type v, float scale) {
  if constexpr (KV == B200PA_KV_AUTO) {
    return Scalar<ST>::to_float(v);
  } else {
    return fp8_dequant_faithful<KV, ST>(v, scale);
  }
}

// grid (num_heads, num_seqs, num_partitions or 1); PARTITIONED selects v2 output.
template <int ST, int KV, int BS, bool PARTITIONED>
__global__ void __launch_bounds__(kGenericThreads)
paged_attention_generic_kernel(
    typename Scalar<ST>::type* __restrict__ out,  // v1: [S,H,D]; v2: tmp_out [S,H,P,D]
    float* __restrict__ lse, int64_t lse_seq_stride, int64_t lse_head_stride,  // v1 only, optional
    float* __restrict__ exp_sums, float* __restrict__ max_logits,
    const typename Scalar<ST>::type* __restrict__ q,
    const typename CacheElem<ST, KV>::type* __restrict__ k_cache,
    const typename CacheElem<ST, KV>::type* __restrict__ v_cache, int num_kv_heads,
    int head_size, float scale, const int32_t* __restrict__ block_tables,
    const int32_t* __restrict__ seq_lens, int max_num_blocks_per_seq,
    const float* __restrict__ alibi_slopes, int64_t q_stride, int64_t kv_block_stride,
    int64_t kv_head_stride, const float* __restrict__ k_scale_ptr,
    const float* __restrict__ v_scale_ptr, int max_num_partitions, BlockSparse bsp) {
  using T = typename Scalar<ST>::type;
  using C = typename CacheElem<ST, KV>::type;
  constexpr int X = 16 / (int)sizeof(C);
  constexpr int TG = (kWarp / BS) > 0 ? (kWarp / BS) : 1;  // lanes per token

  const int head = blockIdx.x, seq = blockIdx.y, part = blockIdx.z;
  const int num_heads = gridDim.x;
  const int seq_len = seq_lens[seq];
  const int tok_begin = PARTITIONED ? part * kPartitionSize : 0;
  if (PARTITIONED && tok_begin >= seq_len) return;
  const int tok_end = PARTITIONED ? min(tok_begin + kPartitionSize, seq_len) : seq_len;
  const int blk_begin = tok_begin / BS;
  const int blk_end = (tok_end + BS - 1) / BS;

  const int warp = threadIdx.x / kWarp, lane = threadIdx.x % kWarp;
  const int kv_head = head / (num_heads / num_kv_heads);
  const float slope = alibi_slopes ? alibi_slopes[head] : 0.f;
  float ks = 1.f, vs = 1.f;
  if constexpr (KV != B200PA_KV_AUTO) { ks = *k_scale_ptr; vs = *v_scale_ptr; }

  __shared__ float q_s[kMaxHead];
  __shared__ float p_s[kGenericWarps][BS];
  __shared__ float m_s[kGenericWarps], l_s[kGenericWarps];
  __shared__ float o_s[kGenericWarps][kMaxHead];

  const T* q_ptr = q + (int64_t)seq * q_stride + (int64_t)head * head_size;
  for (int d = threadIdx.x; d < head_size; d += kGenericThreads) q_s[d] = Scalar<ST>::to_float(q_ptr[d]);
  __syncthreads();

  int bs_offset = 0, q_bs_block = 0;
  if (bsp.enabled) {
    q_bs_block = (seq_len - 1) / bsp.block_size;
    if (bsp.head_sliding_step >= 0)
      bs_offset = (bsp.tp_rank * num_heads + head) * bsp.head_sliding_step + 1;
    else
      bs_offset = (bsp.tp_rank * num_kv_heads + kv_head) * (-bsp.head_sliding_step) + 1;
  }

  const int32_t* block_table = block_tables + (int64_t)seq * max_num_blocks_per_seq;
  const int chunks = head_size / X;
  const int tok_in_blk = lane / TG;  // token handled by this lane in the QK phase
  const int sub = lane % TG;

  float m_run = -INFINITY, l_run = 0.f;
  float acc[kMaxRows];
#pragma unroll
  for (int i = 0; i < kMaxRows; ++i) acc[i] = 0.f;

  for (int blk = blk_begin + warp; blk < blk_end; blk += kGenericWarps) {
    if (bsp.enabled) {
      const int kb = blk * BS / bsp.block_size;
      const bool remote = ((kb + bs_offset) % bsp.vert_stride) == 0;
      const bool local = kb > q_bs_block - bsp.local_blocks;
      if (!remote && !local) continue;  // warp-uniform
    }
    const int64_t phys = block_table[blk];
    const C* kp = k_cache + phys * kv_block_stride + (int64_t)kv_head * kv_head_stride;
    const C* vp = v_cache + phys * kv_block_stride + (int64_t)kv_head * kv_head_stride;

    // ---- q.k for token `tok_in_blk`: lanes of a group split the 16-byte chunks
    float dot = 0.f;
    if (tok_in_blk < BS) {
      for (int c = sub; c < chunks; c += TG) {
        const C* kc = kp + (int64_t)c * BS * X + tok_in_blk * X;
        const uint4 raw = *reinterpret_cast<const uint4*>(kc);
        const C* e = reinterpret_cast<const C*>(&raw);
#pragma unroll
        for (int i = 0; i < X; ++i) dot = fmaf(q_s[c * X + i], cache_to_float<ST, KV>(e[i], ks), dot);
      }
    }
#pragma unroll
    for (int m = TG / 2; m >= 1; m >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, m);
    const int tok = blk * BS + tok_in_blk;
    const bool valid = (tok_in_blk < BS) && (tok < tok_end);
    float logit = scale * dot;
    if (slope != 0.f) logit += slope * (float)(tok - seq_len + 1);

    // ---- online softmax (per warp)
    const float blk_max = warp_max(valid ? logit : -INFINITY);
    const float m_new = fmaxf(m_run, blk_max);  // finite: every processed block has a valid token
    const float alpha = __expf(m_run - m_new);  // m_run == -inf -> 0
    const float p = valid ? __expf(logit - m_new) : 0.f;
    const float p_sum = warp_sum(sub == 0 ? p : 0.f);
    l_run = l_run * alpha + p_sum;
    m_run = m_new;
    if (sub == 0 && tok_in_blk < BS) p_s[warp][tok_in_blk] = p;
    __syncwarp();

    // ---- acc[d] = acc[d]*alpha + sum_t p[t] * V[d][t];  lane owns rows d = lane + 32 i
#pragma unroll
    for (int i = 0; i < kMaxRows; ++i) {
      const int d = lane + kWarp * i;
      if (d < head_size) {
        const C* vr = vp + (int64_t)d * BS;
        float a = acc[i] * alpha;
        constexpr int ROW_BYTES = BS * (int)sizeof(C);
        if constexpr (ROW_BYTES % 16 == 0) {
#pragma unroll
          for (int g = 0; g < ROW_BYTES / 16; ++g) {
            const uint4 raw = reinterpret_cast<const uint4*>(vr)[g];
            const C* e = reinterpret_cast<const C*>(&raw);
#pragma unroll
            for (int j = 0; j < 16 / (int)sizeof(C); ++j) {
              const float pv = p_s[warp][g * (16 / (int)sizeof(C)) + j];
              if (pv != 0.f) a = fmaf(pv, cache_to_float<ST, KV>(e[j], vs), a);
            }
          }
        } else {  // 8-byte rows (block_size 8 with fp8)
          const uint2 raw = *reinterpret_cast<const uint2*>(vr);
          const C* e = reinterpret_cast<const C*>(&raw);
#pragma unroll
          for (int j = 0; j < BS; ++j) {
            const float pv = p_s[warp][j];
            if (pv != 0.f) a = fmaf(pv, cache_to_float<ST, KV>(e[j], vs), a);
          }
        }
        acc[i] = a;
      }
    }
    __syncwarp();
  }

  // ---- merge the warps
  if (lane == 0) { m_s[warp] = m_run; l_s[warp] = l_run; }
  __syncthreads();
  float m_all = -INFINITY;
#pragma unroll
  for (int w = 0; w < kGenericWarps; ++w) m_all = fmaxf(m_all, m_s[w]);
  const float f = (m_run == -INFINITY) ? 0.f : __expf(m_run - m_all);
  float l_all = 0.f;
#pragma unroll
  for (int w = 0; w < kGenericWarps; ++w)
    l_all += (m_s[w] == -INFINITY) ? 0.f : l_s[w] * __expf(m_s[w] - m_all);
#pragma unroll
  for (int i = 0; i < kMaxRows; ++i) {
    const int d = lane + kWarp * i;
    if (d < head_size) o_s[warp][d] = acc[i] * f;
  }
  __syncthreads();
  const float inv = __fdividef(1.f, l_all + 1e-6f);
  T* out_ptr;
  if constexpr (PARTITIONED) {
    out_ptr = out + (((int64_t)seq * num_heads + head) * max_num_partitions + part) * head_size;
    if (threadIdx.x == 0) {
      const int64_t idx = ((int64_t)seq * num_heads + head) * max_num_partitions + part;
      max_logits[idx] = (m_all == -INFINITY) ? -FLT_MAX : m_all;
      exp_sums[idx] = l_all;
    }
  } else {
    out_ptr = out + ((int64_t)seq * num_heads + head) * head_size;
    if (lse != nullptr && threadIdx.x == 0)
      lse[seq * lse_seq_stride + head * lse_head_stride] =
          (m_all == -INFINITY) ? -INFINITY : m_all + logf(l_all);
  }
  for (int d = threadIdx.x; d < head_size; d += kGenericThreads) {
    float o = 0.f;
#pragma unroll
    for (int w = 0; w < kGenericWarps; ++w) o += o_s[w][d];
    out_ptr[d] = Scalar<ST>::from_float(o * inv);
  }
}

// ---------------------------------------------------------------------------
// Split-K reduce (semantics of paged_attention_v2_reduce_kernel,
// attention_kernels.cuh:560-664): w_j = exp_sums[j] * exp(max_j - max),
// out = sum_j tmp_out[j] * w_j / (sum_j w_j + 1e-6).  Slots with exp_sum == 0
// were merged into an earlier slot by the fast kernel and are skipped.
// grid (num_heads, num_seqs), 128 threads.
// ---------------------------------------------------------------------------
template <int ST>
__global__ void __launch_bounds__(kGenericThreads)
paged_attention_reduce_kernel(typename Scalar<ST>::type* __restrict__ out,
                              const float* __restrict__ exp_sums,
                              const float* __restrict__ max_logits,
                              const typename Scalar<ST>::type* __restrict__ tmp_out,
                              const int32_t* __restrict__ seq_lens, int head_size,
                              int max_num_partitions) {
  using T = typename Scalar<ST>::type;
  const int num_heads = gridDim.x, head = blockIdx.x, seq = blockIdx.y;
  const int seq_len = seq_lens[seq];
  const int np = (seq_len + kPartitionSize - 1) / kPartitionSize;
  const int64_t base = ((int64_t)seq * num_heads + head) * max_num_partitions;
  T* out_ptr = out + ((int64_t)seq * num_heads + head) * head_size;
  const T* tmp_ptr = tmp_out + base * head_size;
  if (np == 1) {  // reference shortcut :577-589
    for (int d = threadIdx.x; d < head_size; d += blockDim.x) out_ptr[d] = tmp_ptr[d];
    return;
  }
  extern __shared__ float w_s[];  // [np]
  __shared__ float red[2 * kGenericWarps];
  const int warp = threadIdx.x / kWarp, lane = threadIdx.x % kWarp;

  float mx = -FLT_MAX;
  for (int j = threadIdx.x; j < np; j += blockDim.x)
    if (exp_sums[base + j] != 0.f) mx = fmaxf(mx, max_logits[base + j]);
  mx = warp_max(mx);
  if (lane == 0) red[warp] = mx;
  __syncthreads();
  mx = red[0];
#pragma unroll
  for (int w = 1; w < kGenericWarps; ++w) mx = fmaxf(mx, red[w]);

  float sum = 0.f;
  for (int j = threadIdx.x; j < np; j += blockDim.x) {
    const float l = exp_sums[base + j];
    const float w = (l != 0.f) ? l * expf(max_logits[base + j] - mx) : 0.f;
    w_s[j] = w;
    sum += w;
  }
  sum = warp_sum(sum);
  if (lane == 0) red[kGenericWarps + warp] = sum;
  __syncthreads();
  sum = 0.f;
#pragma unroll
  for (int w = 0; w < kGenericWarps; ++w) sum += red[kGenericWarps + w];
  const float inv = __fdividef(1.f, sum + 1e-6f);

  for (int d = threadIdx.x; d < head_size; d += blockDim.x) {
    float acc = 0.f;
    for (int j = 0; j < np; ++j) {
      const float w = w_s[j];
      if (w != 0.f) acc += Scalar<ST>::to_float(tmp_ptr[(int64_t)j * head_size + d]) * w * inv;
    }
    out_ptr[d] = Scalar<ST>::from_float(acc);
  }
}

template <int ST, int KV, int BS>
int launch_generic_t(void* out, float* lse, int64_t lse_seq_stride, int64_t lse_head_stride,
                     float* exp_sums, float* max_logits, void* tmp_out,
                     const void* query, const void* key_cache, const void* value_cache,
                     int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
                     const int32_t* block_tables, int max_num_blocks_per_seq,
                     const int32_t* seq_lens, int max_seq_len, int max_num_partitions,
                     const float* alibi_slopes, int64_t q_stride, int64_t kv_block_stride,
                     int64_t kv_head_stride, const float* k_scale, const float* v_scale,
                     BlockSparse bsp, bool partitioned, cudaStream_t stream) {
  using T = typename Scalar<ST>::type;
  using C = typename CacheElem<ST, KV>::type;
  dim3 block(kGenericThreads);
  if (!partitioned) {
    dim3 grid(num_heads, num_seqs, 1);
    paged_attention_generic_kernel<ST, KV, BS, false><<<grid, block, 0, stream>>>(
        (T*)out, lse, lse_seq_stride, lse_head_stride, nullptr, nullptr, (const T*)query, (const C*)key_cache, (const C*)value_cache,
        num_kv_heads, head_size, scale, block_tables, seq_lens, max_num_blocks_per_seq,
        alibi_slopes, q_stride, kv_block_stride, kv_head_stride, k_scale, v_scale, 1, bsp);
    count_launch();
    return check_launch("paged_attention_v1 (generic)");
  }
  const int parts = (max_seq_len + kPartitionSize - 1) / kPartitionSize;
  dim3 grid(num_heads, num_seqs, parts > 0 ? parts : 1);
  paged_attention_generic_kernel<ST, KV, BS, true><<<grid, block, 0, stream>>>(
      (T*)tmp_out, nullptr, 0, 0, exp_sums, max_logits, (const T*)query, (const C*)key_cache,
      (const C*)value_cache, num_kv_heads, head_size, scale, block_tables, seq_lens,
      max_num_blocks_per_seq, alibi_slopes, q_stride, kv_block_stride, kv_head_stride, k_scale,
      v_scale, max_num_partitions, bsp);
  count_launch();
  int rc = check_launch("paged_attention_v2 (generic)");
  if (rc) return rc;
  return B200PA_OK;
}

}  // namespace

int launch_reduce(void* out, const float* exp_sums, const float* max_logits, const void* tmp_out,
                  const int32_t* seq_lens, int num_seqs, int num_heads, int head_size,
                  int max_num_partitions, int scalar_type, cudaStream_t stream) {
  dim3 grid(num_heads, num_seqs), block(kGenericThreads);
  const size_t smem = sizeof(float) * (size_t)max_num_partitions;
  if (scalar_type == B200PA_BF16)
    paged_attention_reduce_kernel<B200PA_BF16><<<grid, block, smem, stream>>>(
        (__nv_bfloat16*)out, exp_sums, max_logits, (const __nv_bfloat16*)tmp_out, seq_lens,
        head_size, max_num_partitions);
  else if (scalar_type == B200PA_F16)
    paged_attention_reduce_kernel<B200PA_F16><<<grid, block, smem, stream>>>(
        (__half*)out, exp_sums, max_logits, (const __half*)tmp_out, seq_lens, head_size,
        max_num_partitions);
  else
    paged_attention_reduce_kernel<B200PA_F32><<<grid, block, smem, stream>>>(
        (float*)out, exp_sums, max_logits, (const float*)tmp_out, seq_lens, head_size,
        max_num_partitions);
  count_launch();
  return check_launch("paged_attention_v2 reduce");
}

int launch_generic(void* out, float* lse, int64_t lse_seq_stride, int64_t lse_head_stride,
                   float* exp_sums, float* max_logits, void* tmp_out, const void* query, const void* key_cache, const void* value_cache,
                   int num_seqs, int num_heads, int num_kv_heads, int head_size, int block_size,
                   float scale, const int32_t* block_tables, int max_num_blocks_per_seq,
                   const int32_t* seq_lens, int max_seq_len, int max_num_partitions,
                   const float* alibi_slopes, int64_t q_stride, int64_t kv_block_stride,
                   int64_t kv_head_stride, int scalar_type, int kv_dtype, const float* k_scale,
                   const float* v_scale, BlockSparse bsp, bool partitioned,
                   cudaStream_t stream) {
#define GEN_ARGS                                                                               \
  out, lse, lse_seq_stride, lse_head_stride, exp_sums, max_logits, tmp_out, query, key_cache, \
      value_cache, num_seqs, num_heads, num_kv_heads,                                          \
      head_size, scale, block_tables, max_num_blocks_per_seq, seq_lens,          \
      max_seq_len, max_num_partitions, alibi_slopes, q_stride, kv_block_stride,                \
      kv_head_stride, k_scale, v_scale, bsp, partitioned, stream
#define GEN_BS(ST, KV)                                                                         \
  switch (block_size) {                                                                        \
    case 8: return launch_generic_t<ST, KV, 8>(GEN_ARGS);                                      \
    case 16: return launch_generic_t<ST, KV, 16>(GEN_ARGS);                                    \
    case 32: return launch_generic_t<ST, KV, 32>(GEN_ARGS);                                    \
    default: return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported block size: %d", block_size); \
  }
#define GEN_KV(ST)                                                                             \
  if (kv_dtype == B200PA_KV_AUTO) { GEN_BS(ST, B200PA_KV_AUTO) }                               \
  else if (kv_dtype == B200PA_KV_FP8_E4M3) { GEN_BS(ST, B200PA_KV_FP8_E4M3) }                  \
  else if (kv_dtype == B200PA_KV_FP8_E5M2) { GEN_BS(ST, B200PA_KV_FP8_E5M2) }                  \
  else return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported data type of kv cache: %d", kv_dtype);
  if (scalar_type == B200PA_BF16) { GEN_KV(B200PA_BF16) }
  else if (scalar_type == B200PA_F16) { GEN_KV(B200PA_F16) }
  else if (scalar_type == B200PA_F32) { GEN_KV(B200PA_F32) }
  return set_error(B200PA_ERR_UNSUPPORTED, "Unsupported input type of kv cache: %d", scalar_type);
#undef GEN_KV
#undef GEN_BS
#undef GEN_ARGS
}

}  // namespace b200pa
