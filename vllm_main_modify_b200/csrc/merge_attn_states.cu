// Log-sum-exp merges of partial attention results (sm_100a).
//
//  * merge_attn_states_kernel -- the job of the reference's merge_attn_states
//    (csrc/attention/merge_attn_states.cu:14-89, launcher :131-170, schema
//    csrc/torch_bindings.cpp:80-87): combine a "prefix" and a "suffix" partial
//    attention output with their log-sum-exps (section 2.2 of arXiv 2501.01005),
//        m = max(lse_p, lse_s);  w_p = e^(lse_p - m), w_s = e^(lse_s - m)
//        out = (w_p out_p + w_s out_s) / (w_p + w_s);  lse = log(w_p + w_s) + m
//    with +-inf lse treated as -inf (:45-46).  LSE tensors are [num_heads, num_tokens].
//  * correct_attn_out_kernel -- the job of the reference's Triton
//    _correct_attn_cp_out_kernel (vllm/attention/ops/common.py:9-68): decode context
//    parallelism keeps a slice of every sequence's KV on each rank; after the LSEs
//    of all ranks are all-gathered, each rank rescales its partial output by
//    e^(lse_rank - lse_total) (in place) so that a plain sum over ranks (reduce-scatter)
//    yields the exact attention output.  NaN / +inf LSEs count as -inf (:41, :62-64).
//
// Both are pure streaming kernels (HBM-bound, 16-byte accesses, no reuse).
#include "common.cuh"

namespace b200pa {

namespace {

constexpr int kMergeThreads = 128;

__device__ __forceinline__ float lse_clean_inf(float v) { return isinf(v) ? -INFINITY : v; }

// One thread per 16-byte pack of one (token, head) row, U rows' packs per thread with all loads
// issued before the first use (bytes in flight), 32-bit index arithmetic whenever the problem
// allows (IdxT), streaming (evict-first) accesses: nothing here is read twice.
template <int ST, typename IdxT, int U>
__global__ void __launch_bounds__(kMergeThreads)
merge_attn_states_kernel(typename Scalar<ST>::type* __restrict__ output,
                         float* __restrict__ output_lse,
                         const typename Scalar<ST>::type* __restrict__ prefix_output,
                         const float* __restrict__ prefix_lse,
                         const typename Scalar<ST>::type* __restrict__ suffix_output,
                         const float* __restrict__ suffix_lse, IdxT num_tokens,
                         IdxT num_heads, IdxT packs_per_head, IdxT total) {
  using T = typename Scalar<ST>::type;
  constexpr int PACK = 16 / (int)sizeof(T);
  const IdxT base = (IdxT)blockIdx.x * (IdxT)(kMergeThreads * U) + threadIdx.x;
  IdxT lse_off[U];
  bool live[U], first[U];
  float p_lse[U], s_lse[U];
  uint4 pp[U], sp[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const IdxT idx = base + (IdxT)(u * kMergeThreads);
    live[u] = idx < total;
    const IdxT row = idx / packs_per_head;  // token * num_heads + head
    first[u] = idx == row * packs_per_head;
    const IdxT token = row / num_heads;
    const IdxT head = row - token * num_heads;
    lse_off[u] = head * num_tokens + token;
    if (live[u]) {
      p_lse[u] = __ldg(prefix_lse + lse_off[u]);
      s_lse[u] = __ldg(suffix_lse + lse_off[u]);
      pp[u] = __ldcs(reinterpret_cast<const uint4*>(prefix_output) + idx);
      sp[u] = __ldcs(reinterpret_cast<const uint4*>(suffix_output) + idx);
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    if (!live[u]) continue;
    const IdxT idx = base + (IdxT)(u * kMergeThreads);
    float pl = lse_clean_inf(p_lse[u]), sl = lse_clean_inf(s_lse[u]);
    const float max_lse = fmaxf(pl, sl);
    pl -= max_lse;
    sl -= max_lse;
    const float p_se = expf(pl), s_se = expf(sl);
    const float out_se = p_se + s_se;
    const float p_scale = p_se / out_se, s_scale = s_se / out_se;
    uint4 op;
    const T* pe = reinterpret_cast<const T*>(&pp[u]);
    const T* se = reinterpret_cast<const T*>(&sp[u]);
    T* oe = reinterpret_cast<T*>(&op);
#pragma unroll
    for (int i = 0; i < PACK; ++i)
      oe[i] = Scalar<ST>::from_float(Scalar<ST>::to_float(pe[i]) * p_scale +
                                     (Scalar<ST>::to_float(se[i]) * s_scale));
    __stcs(reinterpret_cast<uint4*>(output) + idx, op);
    if (output_lse != nullptr && first[u]) output_lse[lse_off[u]] = logf(out_se) + max_lse;
  }
}

__device__ __forceinline__ float lse_clean(float v) {
  return (v != v || v == INFINITY) ? -INFINITY : v;
}

// One warp per (sequence, head): total LSE over ranks, then rescale the row in place.
template <int ST>
__global__ void __launch_bounds__(kMergeThreads)
correct_attn_out_kernel(typename Scalar<ST>::type* __restrict__ out, float* __restrict__ lse_out,
                        const float* __restrict__ lses, int num_ranks, int rank, int num_seqs,
                        int num_heads, int head_size, int64_t out_seq_stride,
                        int64_t out_head_stride, int64_t lses_rank_stride,
                        int64_t lses_seq_stride, int64_t lses_head_stride) {
  using T = typename Scalar<ST>::type;
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (kMergeThreads / 32) + (threadIdx.x >> 5);
  if (warp >= (int64_t)num_seqs * num_heads) return;
  const int64_t b = warp / num_heads;
  const int h = (int)(warp - b * num_heads);
  const float* l0 = lses + b * lses_seq_stride + h * lses_head_stride;
  float mx = -INFINITY;
  for (int n = lane; n < num_ranks; n += 32) mx = fmaxf(mx, lse_clean(l0[n * lses_rank_stride]));
  mx = warp_max(mx);
  float acc = 0.f;
  for (int n = lane; n < num_ranks; n += 32) acc += expf(lse_clean(l0[n * lses_rank_stride]) - mx);
  acc = warp_sum(acc);
  const float lse = logf(acc) + mx;  // all ranks empty: -inf - -inf = NaN, as the reference
  if (lane == 0) lse_out[b * lses_seq_stride + h * lses_head_stride] = lse;
  const float factor = expf(lse_clean(l0[rank * lses_rank_stride] - lse));
  T* row = out + b * out_seq_stride + h * out_head_stride;
  for (int d = lane; d < head_size; d += 32)
    row[d] = Scalar<ST>::from_float(Scalar<ST>::to_float(row[d]) * factor);
}

// 16-byte variant (rows 16-byte aligned, head_size a multiple of the pack).  A CTA owns
// kCorrPacks consecutive 16-byte packs (64 rows at head size 128, 16-bit): every thread first
// issues the loads of its kCorrU packs, then one thread per ROW computes that row's total LSE
// and factor (the ranks' LSEs 8 loads at a time) into shared memory, then the packs are scaled
// and stored.  (One thread per pack each recomputing the row's log-sum-exp was issue-bound:
// ~350 instructions per 32 bytes moved, 0.39 of the HBM peak in tools/opbench.py.)
constexpr int kCorrThreads = 256, kCorrU = 4, kCorrPacks = kCorrThreads * kCorrU;

template <int ST>
__global__ void __launch_bounds__(kCorrThreads)
correct_attn_out_vec_kernel(typename Scalar<ST>::type* __restrict__ out,
                            float* __restrict__ lse_out, const float* __restrict__ lses,
                            int num_ranks, int rank, unsigned total, unsigned num_heads,
                            unsigned packs_per_head, int64_t out_seq_stride,
                            int64_t out_head_stride, int64_t lses_rank_stride,
                            int64_t lses_seq_stride, int64_t lses_head_stride) {
  using T = typename Scalar<ST>::type;
  constexpr int PACK = 16 / (int)sizeof(T);
  __shared__ float fac_s[kCorrPacks + 1];
  const unsigned p0 = blockIdx.x * kCorrPacks;
  const unsigned p1 = min(p0 + (unsigned)kCorrPacks, total);
  const unsigned row0 = p0 / packs_per_head, row1 = (p1 - 1) / packs_per_head;
  uint4* ptr[kCorrU];
  uint4 v[kCorrU];
  unsigned lrow[kCorrU];
#pragma unroll
  for (int u = 0; u < kCorrU; ++u) {
    const unsigned idx = p0 + u * kCorrThreads + threadIdx.x;
    const unsigned row = idx / packs_per_head;
    const unsigned pack = idx - row * packs_per_head;
    const unsigned b = row / num_heads;
    const unsigned h = row - b * num_heads;
    lrow[u] = row - row0;
    ptr[u] = reinterpret_cast<uint4*>(out + b * out_seq_stride + h * out_head_stride) + pack;
    if (idx < p1) v[u] = __ldcs(ptr[u]);
  }
  for (unsigned i = threadIdx.x; i <= row1 - row0; i += kCorrThreads) {
    const unsigned row = row0 + i;
    const unsigned b = row / num_heads;
    const unsigned h = row - b * num_heads;
    const float* l0 = lses + b * lses_seq_stride + h * lses_head_stride;
    const float mine = lse_clean(__ldg(l0 + rank * lses_rank_stride));
    constexpr int CH = 8;
    float mx = -INFINITY, acc = 0.f;
    for (int n0 = 0; n0 < num_ranks; n0 += CH) {
      float l[CH];
#pragma unroll
      for (int j = 0; j < CH; ++j)
        l[j] = (n0 + j < num_ranks) ? lse_clean(__ldg(l0 + (n0 + j) * lses_rank_stride)) : -INFINITY;
      float nm = mx;
#pragma unroll
      for (int j = 0; j < CH; ++j) nm = fmaxf(nm, l[j]);
      if (nm == -INFINITY) continue;   // nothing but empty ranks so far
      float sum = expf(l[0] - nm);
#pragma unroll
      for (int j = 1; j < CH; ++j)
        if (n0 + j < num_ranks) sum += expf(l[j] - nm);
      acc = acc * expf(mx - nm) + sum;
      mx = nm;
    }
    // every rank empty: -inf - -inf = NaN in the reference (common.py:44-50)
    const float lse = mx == -INFINITY ? __int_as_float(0x7fc00000) : logf(acc) + mx;
    if (row * packs_per_head >= p0)   // the CTA that holds the row's first pack
      lse_out[b * lses_seq_stride + h * lses_head_stride] = lse;
    fac_s[i] = expf(lse_clean(mine - lse));
  }
  __syncthreads();
#pragma unroll
  for (int u = 0; u < kCorrU; ++u) {
    if (p0 + u * kCorrThreads + threadIdx.x >= p1) continue;
    const float factor = fac_s[lrow[u]];
    uint4 o;
    const T* ve = reinterpret_cast<const T*>(&v[u]);
    T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int i = 0; i < PACK; ++i)
      oe[i] = Scalar<ST>::from_float(Scalar<ST>::to_float(ve[i]) * factor);
    __stcs(ptr[u], o);
  }
}

}  // namespace

int launch_merge_attn_states(void* output, float* output_lse, const void* prefix_output,
                             const float* prefix_lse, const void* suffix_output,
                             const float* suffix_lse, int64_t num_tokens, int num_heads,
                             int head_size, int scalar_type, cudaStream_t stream) {
  const int pack = scalar_type == B200PA_F32 ? 4 : 8;
  const int pph = head_size / pack;
  const int64_t total = num_tokens * num_heads * pph;
  if (total == 0) return 0;
  constexpr int U = 2;
  const int64_t blocks = (total + kMergeThreads * U - 1) / (kMergeThreads * U);
  if (blocks > 0x7fffffffLL) return set_error(B200PA_ERR_UNSUPPORTED, "merge_attn_states: too large");
  const dim3 grid((unsigned)blocks), block(kMergeThreads);
  // 32-bit indices when every offset (incl. the rounded-up tail) fits
  const bool small = total + kMergeThreads * U < (1LL << 32);
#define MERGE_CALL(ST)                                                                      \
  if (small)                                                                                \
    merge_attn_states_kernel<ST, uint32_t, U><<<grid, block, 0, stream>>>(                  \
        (typename Scalar<ST>::type*)output, output_lse,                                     \
        (const typename Scalar<ST>::type*)prefix_output, prefix_lse,                        \
        (const typename Scalar<ST>::type*)suffix_output, suffix_lse, (uint32_t)num_tokens,  \
        (uint32_t)num_heads, (uint32_t)pph, (uint32_t)total);                               \
  else                                                                                      \
    merge_attn_states_kernel<ST, int64_t, U><<<grid, block, 0, stream>>>(                   \
        (typename Scalar<ST>::type*)output, output_lse,                                     \
        (const typename Scalar<ST>::type*)prefix_output, prefix_lse,                        \
        (const typename Scalar<ST>::type*)suffix_output, suffix_lse, (int64_t)num_tokens,   \
        (int64_t)num_heads, (int64_t)pph, (int64_t)total)
  if (scalar_type == B200PA_BF16) { MERGE_CALL(B200PA_BF16); }
  else if (scalar_type == B200PA_F16) { MERGE_CALL(B200PA_F16); }
  else { MERGE_CALL(B200PA_F32); }
#undef MERGE_CALL
  count_launch();
  return check_launch("merge_attn_states");
}

int launch_correct_attn_out(void* out, float* lse_out, const float* lses, int num_ranks, int rank,
                            int num_seqs, int num_heads, int head_size, int64_t out_seq_stride,
                            int64_t out_head_stride, int64_t lses_rank_stride,
                            int64_t lses_seq_stride, int64_t lses_head_stride, int scalar_type,
                            cudaStream_t stream) {
  const int64_t warps = (int64_t)num_seqs * num_heads;
  if (warps == 0) return 0;
  const int pack = scalar_type == B200PA_F32 ? 4 : 8;
  const int64_t vec_total = warps * (head_size / pack);
  if (head_size % pack == 0 && out_seq_stride % pack == 0 && out_head_stride % pack == 0 &&
      ((uintptr_t)out & 15) == 0 && vec_total < (1LL << 32) - kCorrPacks) {
    const dim3 vgrid((unsigned)((vec_total + kCorrPacks - 1) / kCorrPacks));
#define CORRECT_VEC_CALL(ST)                                                                 \
  correct_attn_out_vec_kernel<ST><<<vgrid, kCorrThreads, 0, stream>>>(                      \
      (typename Scalar<ST>::type*)out, lse_out, lses, num_ranks, rank, (unsigned)vec_total,  \
      (unsigned)num_heads, (unsigned)(head_size / pack), out_seq_stride, out_head_stride,    \
      lses_rank_stride, lses_seq_stride, lses_head_stride)
    if (scalar_type == B200PA_BF16) CORRECT_VEC_CALL(B200PA_BF16);
    else if (scalar_type == B200PA_F16) CORRECT_VEC_CALL(B200PA_F16);
    else CORRECT_VEC_CALL(B200PA_F32);
#undef CORRECT_VEC_CALL
    count_launch();
    return check_launch("correct_attn_out");
  }
  const int wpb = kMergeThreads / 32;
  const dim3 grid((unsigned)((warps + wpb - 1) / wpb)), block(kMergeThreads);
#define CORRECT_CALL(ST)                                                                    \
  correct_attn_out_kernel<ST><<<grid, block, 0, stream>>>(                                  \
      (typename Scalar<ST>::type*)out, lse_out, lses, num_ranks, rank, num_seqs, num_heads, \
      head_size, out_seq_stride, out_head_stride, lses_rank_stride, lses_seq_stride,        \
      lses_head_stride)
  if (scalar_type == B200PA_BF16) CORRECT_CALL(B200PA_BF16);
  else if (scalar_type == B200PA_F16) CORRECT_CALL(B200PA_F16);
  else CORRECT_CALL(B200PA_F32);
#undef CORRECT_CALL
  count_launch();
  return check_launch("correct_attn_out");
}

}  // namespace b200pa
