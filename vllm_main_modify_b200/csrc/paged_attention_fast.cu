// TMA-staged, tensor-core paged-decode attention for sm_100a.
//
// Replaces the per-sequence decode kernel of the reference
// (csrc/attention/attention_kernels.cuh:85-491, launched from
// paged_attention_v1.cu:46-125 / paged_attention_v2.cu:47-126) for the shapes of
// the BASELINE configs: head_size 64/128, block_size 16, bf16/fp16 query,
// kv cache "auto" (16-bit) or fp8_e4m3.  Design:
//
//  * One CTA per (token chunk, kv head, sequence).  All G = num_heads/num_kv_heads
//    query heads of the group share every K/V page load (the reference re-reads
//    each page G times, once per query-head CTA).
//  * Each warp owns a private ring of STAGES page buffers in shared memory and
//    walks pages warp, warp+NW, ... of the chunk.  Lane 0 issues one
//    `cp.async.bulk` (TMA, 1-D) per K page and per V page -- the x-packed K
//    sub-block [D/x][16][x] and the V sub-block [D][16] of one (page, kv head)
//    are each one contiguous 2 KiB (fp8) / 4 KiB (16-bit) run -- completing on
//    a per-stage mbarrier.  No producer warp, no cross-warp synchronisation in
//    the main loop; block-table entries are prefetched 32 at a time per warp.
//  * QK^T and PV run on `mma.sync.m16n8k16` with the query group padded to the
//    16-row M tile (fp32 accumulate).  16-bit pages feed the B fragments straight
//    from shared memory with `ldmatrix` (the x-packed K layout is exactly a stack
//    of 8x8 b16 matrices).  fp8 pages are converted in registers with
//    `cvt.rn.f16x2.e4m3x2` (exact) and multiplied in fp16; k_scale is folded into
//    the softmax scale and v_scale into the output scale.
//  * Online softmax in the log2 domain with lazy rescaling: the subtracted
//    reference max only moves when the running max grew by more than 2^8.
//  * Split-K: a chunk is a whole number of 512-token partitions; the partial is
//    written to the first partition slot of the chunk, the other slots of the
//    chunk get exp_sum = 0, and paged_attention_reduce_kernel merges.
#include "common.cuh"

namespace b200pa {

namespace {

constexpr int kBS = 16;                      // tokens per page (block_size)
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// ------------------------------------------------------------------ PTX bits
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
// 1-D TMA bulk copy global -> shared, completion on an mbarrier.
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void* src, uint32_t bytes,
                                            uint32_t bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
template <bool BF16>
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                         uint32_t b1) {
  if constexpr (BF16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// pack (lo, hi) floats into a 16-bit pair
template <bool BF16>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  uint32_t r;
  if constexpr (BF16) {
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  } else {
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  }
  return r;
}
// e4m3 x4 (one 32-bit word) -> two f16x2 words: lo = bytes 0,1 ; hi = bytes 2,3
__device__ __forceinline__ void e4m3x4_to_f16x2x2(uint32_t w, uint32_t& lo, uint32_t& hi) {
  asm("{\n"
      ".reg .b16 l, h;\n"
      "mov.b32 {l, h}, %2;\n"
      "cvt.rn.f16x2.e4m3x2 %0, l;\n"
      "cvt.rn.f16x2.e4m3x2 %1, h;\n"
      "}\n"
      : "=r"(lo), "=r"(hi)
      : "r"(w));
}

struct FastParams {
  void* out;           // v1: out [S,H,D]; v2: tmp_out [S,H,P,D]
  float* exp_sums;     // v2 only
  float* max_logits;   // v2 only
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const float* alibi_slopes;
  const float* k_scale;
  const float* v_scale;
  int64_t q_stride, kv_block_stride, kv_head_stride;  // elements
  int num_heads, num_kv_heads, group;                  // group = num_heads / num_kv_heads
  int m_tiles;                                         // ceil(group / 16)
  int max_num_blocks_per_seq, max_num_partitions;
  int parts_per_chunk;                                 // v2: 512-token partitions per CTA
  float scale;
};

// token of the page that S column pair `half` (0: s0, 1: s1) of quad-lane q covers,
// after the fp8 swap (see the kernel): 16-bit pages use the identity column order.
template <bool FP8> __device__ __forceinline__ int tok_of(int q, int half) {
  return FP8 ? (4 * q + 2 * half) : (2 * q + 8 * half);
}

// ST: B200PA_F16 | B200PA_BF16 (query/out type).  FP8: kv cache is e4m3 bytes.
template <int ST, bool FP8, int D, int NW, int STAGES, bool PARTITIONED>
__global__ void __launch_bounds__(NW * 32)
paged_attention_fast_kernel(const FastParams p) {
  using T = typename Scalar<ST>::type;
  constexpr bool MMA_BF16 = (!FP8) && (ST == B200PA_BF16);  // fp8 pages compute in f16
  constexpr int CES = FP8 ? 1 : 2;                           // cache element bytes
  constexpr int X = 16 / CES;
  constexpr int PAGE_BYTES = D * kBS * CES;                  // one K (or V) page of one kv head
  constexpr int STAGE_BYTES = 2 * PAGE_BYTES;
  constexpr int KSTEPS = D / 16;                             // QK k-steps
  constexpr int NT = D / 8;                                  // PV n-tiles
  constexpr int RING_BYTES = STAGES * STAGE_BYTES;
  static_assert(RING_BYTES >= 16 * D * 4 + 3 * 16 * 4, "ring too small for the merge buffer");
  // probabilities are exp2(t - msub) with t - msub <= CAP; msub trails the max by OFF
  constexpr float OFF = MMA_BF16 ? 0.f : 6.f;
  constexpr float CAP = MMA_BF16 ? 8.f : 14.f;

  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[NW * STAGES];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = lane >> 2, q4 = lane & 3;
  const int seq = blockIdx.z;
  const int kv_head = blockIdx.y / p.m_tiles;
  const int m_tile = blockIdx.y - kv_head * p.m_tiles;
  const int chunk = blockIdx.x;
  const int seq_len = p.seq_lens[seq];

  int tok_begin = 0, tok_end = seq_len;
  if constexpr (PARTITIONED) {
    tok_begin = chunk * p.parts_per_chunk * kPartitionSize;
    if (tok_begin >= seq_len) return;
    tok_end = min(seq_len, tok_begin + p.parts_per_chunk * kPartitionSize);
  }
  const int page_begin = tok_begin / kBS;
  const int n_pages = (tok_end - tok_begin + kBS - 1) / kBS;
  const int rows_valid = min(16, p.group - m_tile * 16);     // query heads in this M tile
  const int head0 = kv_head * p.group + m_tile * 16;

  if (n_pages <= 0) {  // empty sequence (v1 only): the reference yields zeros
    for (int i = threadIdx.x; i < rows_valid * D; i += NW * 32) {
      const int row = i / D, d = i - row * D;
      reinterpret_cast<T*>(p.out)[((int64_t)seq * p.num_heads + head0 + row) * D + d] =
          Scalar<ST>::from_float(0.f);
    }
    return;
  }

  // ---- per-warp ring + barriers
  uint8_t* ring = smem + (size_t)warp * RING_BYTES;
  const uint32_t ring_u32 = smem_u32(ring);
  const uint32_t bar0 = smem_u32(&bars[warp * STAGES]);
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(bar0 + 8 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();

  const int my_pages = (warp < n_pages) ? (n_pages - warp + NW - 1) / NW : 0;
  const int32_t* bt = p.block_tables + (int64_t)seq * p.max_num_blocks_per_seq + page_begin;
  const uint8_t* kc_base = reinterpret_cast<const uint8_t*>(p.k_cache) +
                           (int64_t)kv_head * p.kv_head_stride * CES;
  const uint8_t* vc_base = reinterpret_cast<const uint8_t*>(p.v_cache) +
                           (int64_t)kv_head * p.kv_head_stride * CES;
  const int64_t blk_stride_bytes = p.kv_block_stride * CES;
  const uint64_t policy = l2_evict_first_policy();

  int bt_reg = 0;  // block-table entries of my pages [32*b, 32*b+32), one per lane
  auto issue = [&](int i) {
    if ((i & 31) == 0) {
      const int idx = i + lane;
      bt_reg = (idx < my_pages) ? bt[warp + NW * idx] : 0;
    }
    const int phys = __shfl_sync(0xffffffffu, bt_reg, i & 31);
    if (lane == 0) {
      const int s = i % STAGES;
      const uint32_t bar = bar0 + 8 * s;
      const uint32_t dst = ring_u32 + s * STAGE_BYTES;
      mbar_expect_tx(bar, STAGE_BYTES);
      tma_load_1d(dst, kc_base + (int64_t)phys * blk_stride_bytes, PAGE_BYTES, bar, policy);
      tma_load_1d(dst + PAGE_BYTES, vc_base + (int64_t)phys * blk_stride_bytes, PAGE_BYTES, bar,
                  policy);
    }
  };
  // start the pipeline before touching Q
  {
    const int pre = min(STAGES, my_pages);
    for (int i = 0; i < pre; ++i) issue(i);
  }

  // ---- Q fragments (A operand), rows r and r+8 of the M tile
  uint32_t qa[KSTEPS][4];
  {
    const T* qbase = reinterpret_cast<const T*>(p.q) + (int64_t)seq * p.q_stride;
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int row = r + 8 * (j & 1);
        // 16-bit pages: natural k order.  fp8 pages: k-slots (2q,2q+1 | 2q+8,2q+9) of
        // k-step ks hold head elements 16ks + (4q,4q+1 | 4q+2,4q+3), matching the
        // 4 consecutive bytes a lane gets from an fp8 row via ldmatrix.
        const int d = FP8 ? (16 * ks + 4 * q4 + 2 * (j >> 1)) : (16 * ks + 2 * q4 + 8 * (j >> 1));
        uint32_t v = 0;
        if (row < rows_valid) {
          const T* src = qbase + (int64_t)(head0 + row) * D + d;
          if constexpr (FP8 && ST == B200PA_BF16) {
            const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(src);
            v = pack2<false>(__bfloat162float(b.x), __bfloat162float(b.y));
          } else {
            v = *reinterpret_cast<const uint32_t*>(src);
          }
        }
        qa[ks][j] = v;
      }
    }
  }

  float slope_l2[2] = {0.f, 0.f};
  const bool has_alibi = p.alibi_slopes != nullptr;
  if (has_alibi) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int row = r + 8 * j;
      if (row < rows_valid) slope_l2[j] = p.alibi_slopes[head0 + row] * kLog2e;
    }
  }
  float sl2 = p.scale * kLog2e;
  float out_scale = 1.f;
  if constexpr (FP8) {
    sl2 *= *p.k_scale;
    out_scale = *p.v_scale;
  }

  float o[NT][4];
#pragma unroll
  for (int n = 0; n < NT; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
  float msub[2] = {-INFINITY, -INFINITY};   // subtracted reference (log2 domain)
  float mtrue[2] = {-INFINITY, -INFINITY};  // true running max (log2 domain)
  float lsum[2] = {0.f, 0.f};               // per-lane partial of sum exp2(t - msub)

  // ldmatrix lane addressing
  const int mat = lane >> 3, mi = lane & 7;
  uint32_t k_lane_off, v_lane_off;
  if constexpr (FP8) {
    // x4 = {chunk c, tile0 | chunk c, tile1 | chunk c+1, tile0 | chunk c+1, tile1};
    // row i of tile 0 is token sigma0(i) = 4(i/2) + (i%2) + (i >= 4 ? 2 : 0), tile 1 is
    // sigma0 ^ 2: both hit 8 distinct 16-byte bank groups, and quad-lane q ends up with
    // tokens 4q..4q+3 -- the 4 bytes it later gets from a V row.
    const int sig0 = 4 * (mi >> 1) + (mi & 1) + ((mi >= 4) ? 2 : 0);
    const int tok = (mat & 1) ? (sig0 ^ 2) : sig0;
    k_lane_off = ((mat >> 1) * kBS + tok) * 16;
    v_lane_off = lane * 16;  // 32 consecutive V rows per x4
  } else {
    // x4 = {chunk 2ks, tok 0-7 | chunk 2ks+1, tok 0-7 | chunk 2ks, tok 8-15 | chunk 2ks+1, tok 8-15}
    k_lane_off = ((mat & 1) * kBS + (mat >> 1) * 8 + mi) * 16;
    // x4 = {d 0-7, tok 0-7 | d 0-7, tok 8-15 | d 8-15, tok 0-7 | d 8-15, tok 8-15}
    v_lane_off = ((mat >> 1) * 8 + mi) * 32 + (mat & 1) * 16;
  }

  for (int i = 0; i < my_pages; ++i) {
    const int s = i % STAGES;
    const uint32_t kbase = ring_u32 + s * STAGE_BYTES;
    const uint32_t vbase = kbase + PAGE_BYTES;
    const int page = warp + NW * i;                       // page index within the chunk
    const int page_tok0 = tok_begin + page * kBS;         // first token of the page
    const int n_valid = min(kBS, tok_end - page_tok0);    // tokens of the page inside the context
    mbar_wait(bar0 + 8 * s, (i / STAGES) & 1);

    if (n_valid < kBS) {
      // tail page: zero the V columns past the context (they may hold NaN), as the
      // reference does at attention_kernels.cuh:415-425
      uint8_t* vpage = ring + s * STAGE_BYTES + PAGE_BYTES;
      for (int d = lane; d < D; d += 32)
        for (int t = n_valid; t < kBS; ++t) {
          if constexpr (FP8) vpage[d * kBS + t] = 0;
          else reinterpret_cast<uint16_t*>(vpage)[d * kBS + t] = 0;
        }
      __syncwarp();
    }

    // ---- S = Q K^T : 16 x 16 tile, two column tiles, two independent k chains each
    float sa[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    float sb[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    if constexpr (FP8) {
#pragma unroll
      for (int c = 0; c < KSTEPS; c += 2) {
        uint32_t w[4];
        ldmatrix_x4(w, kbase + c * (kBS * 16) + k_lane_off);
        uint32_t b0, b1;
        e4m3x4_to_f16x2x2(w[0], b0, b1); mma16816<false>(sa[0], qa[c], b0, b1);
        e4m3x4_to_f16x2x2(w[1], b0, b1); mma16816<false>(sa[1], qa[c], b0, b1);
        e4m3x4_to_f16x2x2(w[2], b0, b1); mma16816<false>(sb[0], qa[c + 1], b0, b1);
        e4m3x4_to_f16x2x2(w[3], b0, b1); mma16816<false>(sb[1], qa[c + 1], b0, b1);
      }
    } else {
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        uint32_t w[4];
        ldmatrix_x4(w, kbase + ks * (2 * kBS * 16) + k_lane_off);
        if (ks & 1) {
          mma16816<MMA_BF16>(sb[0], qa[ks], w[0], w[1]);
          mma16816<MMA_BF16>(sb[1], qa[ks], w[2], w[3]);
        } else {
          mma16816<MMA_BF16>(sa[0], qa[ks], w[0], w[1]);
          mma16816<MMA_BF16>(sa[1], qa[ks], w[2], w[3]);
        }
      }
    }
    // t[half][j]: half 0/1 = column pair, j: 0,1 row r ; 2,3 row r+8
    float t[2][4];
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
      for (int j = 0; j < 4; ++j) t[h][j] = (sa[h][j] + sb[h][j]) * sl2;
    if constexpr (FP8) {
      // quad-lanes 2,3 hold their column pairs swapped (sigma above): restore so that
      // half 0 = tokens 4q,4q+1 and half 1 = tokens 4q+2,4q+3 for every lane
      if (q4 >= 2) {
#pragma unroll
        for (int j = 0; j < 4; ++j) { const float x = t[0][j]; t[0][j] = t[1][j]; t[1][j] = x; }
      }
    }
    if (has_alibi) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int tok = page_tok0 + tok_of<FP8>(q4, h) + (j & 1);
          t[h][j] += slope_l2[j >> 1] * (float)(tok - seq_len + 1);
        }
    }
    if (n_valid < kBS) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (tok_of<FP8>(q4, h) + (j & 1) >= n_valid) t[h][j] = -INFINITY;
    }

    // ---- online softmax, rows r (j=0) and r+8 (j=1)
    float alpha[2];
    bool moved = false;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      float mx = fmaxf(fmaxf(t[0][2 * j], t[0][2 * j + 1]), fmaxf(t[1][2 * j], t[1][2 * j + 1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      mtrue[j] = fmaxf(mtrue[j], mx);
      alpha[j] = 1.f;
      if (mx > msub[j] + CAP) {          // includes the first page (msub = -inf)
        const float nm = mx - OFF;
        alpha[j] = ex2(msub[j] - nm);    // ex2(-inf) = 0
        msub[j] = nm;
        moved = true;
      }
    }
    if (__any_sync(0xffffffffu, moved)) {
#pragma unroll
      for (int n = 0; n < NT; ++n) {
        o[n][0] *= alpha[0]; o[n][1] *= alpha[0];
        o[n][2] *= alpha[1]; o[n][3] *= alpha[1];
      }
      lsum[0] *= alpha[0];
      lsum[1] *= alpha[1];
    }
    uint32_t pa[4];
    {
      float pr[2][4];
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int j = 0; j < 4; ++j) pr[h][j] = ex2(t[h][j] - msub[j >> 1]);
      lsum[0] += (pr[0][0] + pr[0][1]) + (pr[1][0] + pr[1][1]);
      lsum[1] += (pr[0][2] + pr[0][3]) + (pr[1][2] + pr[1][3]);
      // A fragment of P: a0 row r / slots 2q,2q+1 ; a1 row r+8 ; a2,a3 slots 2q+8,2q+9
      pa[0] = pack2<MMA_BF16>(pr[0][0], pr[0][1]);
      pa[1] = pack2<MMA_BF16>(pr[0][2], pr[0][3]);
      pa[2] = pack2<MMA_BF16>(pr[1][0], pr[1][1]);
      pa[3] = pack2<MMA_BF16>(pr[1][2], pr[1][3]);
    }

    // ---- O += P V
    if constexpr (FP8) {
#pragma unroll
      for (int g = 0; g < D / 32; ++g) {
        uint32_t w[4];
        ldmatrix_x4(w, vbase + g * (32 * 16) + v_lane_off);
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          uint32_t b0, b1;
          e4m3x4_to_f16x2x2(w[m], b0, b1);
          mma16816<false>(o[4 * g + m], pa, b0, b1);
        }
      }
    } else {
#pragma unroll
      for (int g = 0; g < D / 16; ++g) {
        uint32_t w[4];
        ldmatrix_x4(w, vbase + g * (16 * 32) + v_lane_off);
        mma16816<MMA_BF16>(o[2 * g], pa, w[0], w[1]);
        mma16816<MMA_BF16>(o[2 * g + 1], pa, w[2], w[3]);
      }
    }

    // ---- refill this stage (all lanes are past their last read of it)
    __syncwarp();
    if (i + STAGES < my_pages) issue(i + STAGES);
  }

  // ---- per-warp partial -> shared (the warp's own, now idle, ring), then merge
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    lsum[j] += __shfl_xor_sync(0xffffffffu, lsum[j], 1);
    lsum[j] += __shfl_xor_sync(0xffffffffu, lsum[j], 2);
  }
  float* wo = reinterpret_cast<float*>(ring);            // [16][D]
  float* wstat = wo + 16 * D;                            // msub[16], mtrue[16], l[16]
#pragma unroll
  for (int n = 0; n < NT; ++n) {
    *reinterpret_cast<float2*>(&wo[r * D + 8 * n + 2 * q4]) = make_float2(o[n][0], o[n][1]);
    *reinterpret_cast<float2*>(&wo[(r + 8) * D + 8 * n + 2 * q4]) = make_float2(o[n][2], o[n][3]);
  }
  if (q4 == 0) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      wstat[r + 8 * j] = msub[j];
      wstat[16 + r + 8 * j] = mtrue[j];
      wstat[32 + r + 8 * j] = lsum[j];
    }
  }
  __syncthreads();

  // every thread: rows_valid * D/2 output pairs spread over the CTA
  for (int e = threadIdx.x; e < rows_valid * (D / 2); e += NW * 32) {
    const int row = e / (D / 2);
    const int d = (e - row * (D / 2)) * 2;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < NW; ++w)
      M = fmaxf(M, reinterpret_cast<const float*>(smem + (size_t)w * RING_BYTES)[16 * D + 16 + row]);
    float L = 0.f, a0 = 0.f, a1 = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const float* pw = reinterpret_cast<const float*>(smem + (size_t)w * RING_BYTES);
      const float ms = pw[16 * D + row];
      const float f = (ms == -INFINITY) ? 0.f : ex2(ms - M);
      L += pw[16 * D + 32 + row] * f;
      const float2 v = *reinterpret_cast<const float2*>(&pw[row * D + d]);
      a0 += v.x * f;
      a1 += v.y * f;
    }
    const float inv = out_scale * __fdividef(1.f, L + 1e-6f);
    const int head = head0 + row;
    T* dst;
    if constexpr (PARTITIONED) {
      const int slot = chunk * p.parts_per_chunk;
      const int64_t idx = ((int64_t)seq * p.num_heads + head) * p.max_num_partitions + slot;
      dst = reinterpret_cast<T*>(p.out) + idx * D + d;
      if (d == 0) {
        p.max_logits[idx] = M * kLn2;
        p.exp_sums[idx] = L;
        // the other partitions of this chunk were merged into `slot`
        const int np = (seq_len + kPartitionSize - 1) / kPartitionSize;
        const int last = min(np, slot + p.parts_per_chunk);
        for (int j = slot + 1; j < last; ++j) {
          p.exp_sums[idx - slot + j] = 0.f;
          p.max_logits[idx - slot + j] = -FLT_MAX;
        }
      }
    } else {
      dst = reinterpret_cast<T*>(p.out) + ((int64_t)seq * p.num_heads + head) * D + d;
    }
    if constexpr (ST == B200PA_BF16) {
      *reinterpret_cast<uint32_t*>(dst) = pack2<true>(a0 * inv, a1 * inv);
    } else {
      *reinterpret_cast<uint32_t*>(dst) = pack2<false>(a0 * inv, a1 * inv);
    }
  }
}

template <bool FP8, int D> struct FastCfg {
  // 4 warps per CTA, two CTAs per SM; ring of 24 KiB per warp
  static constexpr int NW = 4;
  static constexpr int STAGES = (24 * 1024) / (2 * D * kBS * (FP8 ? 1 : 2));
};

template <int ST, bool FP8, int D>
int launch_fast_t(const FastParams& p, int num_seqs, int n_chunks, bool partitioned,
                  cudaStream_t stream) {
  constexpr int NW = FastCfg<FP8, D>::NW;
  constexpr int STAGES = FastCfg<FP8, D>::STAGES;
  constexpr int SMEM = NW * STAGES * 2 * D * kBS * (FP8 ? 1 : 2);
  dim3 grid(n_chunks, p.num_kv_heads * p.m_tiles, num_seqs), block(NW * 32);
  cudaError_t e;
  if (partitioned) {
    auto kern = paged_attention_fast_kernel<ST, FP8, D, NW, STAGES, true>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) return set_error(B200PA_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    kern<<<grid, block, SMEM, stream>>>(p);
  } else {
    auto kern = paged_attention_fast_kernel<ST, FP8, D, NW, STAGES, false>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) return set_error(B200PA_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    kern<<<grid, block, SMEM, stream>>>(p);
  }
  count_launch();
  return check_launch(partitioned ? "paged_attention_v2 (fast)" : "paged_attention_v1 (fast)");
}

inline bool aligned_to(const void* p, uintptr_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

}  // namespace

// Is (shape, dtype, layout) covered by the tensor-core kernel?
bool fast_supported(const void* query, const void* key_cache, const void* value_cache,
                    int num_heads, int num_kv_heads, int head_size, int block_size,
                    int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
                    int scalar_type, int kv_dtype, int blocksparse_vert_stride) {
  if (blocksparse_vert_stride > 1) return false;
  if (block_size != kBS) return false;
  if (head_size != 64 && head_size != 128) return false;
  if (scalar_type != B200PA_BF16 && scalar_type != B200PA_F16) return false;
  if (kv_dtype != B200PA_KV_AUTO && kv_dtype != B200PA_KV_FP8_E4M3) return false;
  if (num_kv_heads <= 0 || num_heads % num_kv_heads != 0) return false;
  const int ces = kv_dtype == B200PA_KV_AUTO ? 2 : 1;
  // TMA bulk copies need 16-byte aligned sources; Q is read as 32-bit pairs
  if (!aligned_to(key_cache, 16) || !aligned_to(value_cache, 16) || !aligned_to(query, 4)) return false;
  if ((kv_block_stride * ces) % 16 != 0 || (kv_head_stride * ces) % 16 != 0) return false;
  if (kv_head_stride < (int64_t)head_size * block_size) return false;
  if (q_stride % 2 != 0) return false;
  return true;
}

int launch_fast(void* out, float* exp_sums, float* max_logits, void* tmp_out, const void* query,
                const void* key_cache, const void* value_cache, int num_seqs, int num_heads,
                int num_kv_heads, int head_size, float scale, const int32_t* block_tables,
                int max_num_blocks_per_seq, const int32_t* seq_lens, int max_seq_len,
                int max_num_partitions, const float* alibi_slopes, int64_t q_stride,
                int64_t kv_block_stride, int64_t kv_head_stride, int scalar_type, int kv_dtype,
                const float* k_scale, const float* v_scale, bool partitioned, int num_sms,
                cudaStream_t stream) {
  FastParams p;
  p.out = partitioned ? tmp_out : out;
  p.exp_sums = exp_sums;
  p.max_logits = max_logits;
  p.q = query;
  p.k_cache = key_cache;
  p.v_cache = value_cache;
  p.block_tables = block_tables;
  p.seq_lens = seq_lens;
  p.alibi_slopes = alibi_slopes;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.q_stride = q_stride;
  p.kv_block_stride = kv_block_stride;
  p.kv_head_stride = kv_head_stride;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.group = num_heads / num_kv_heads;
  p.m_tiles = (p.group + 15) / 16;
  p.max_num_blocks_per_seq = max_num_blocks_per_seq;
  p.max_num_partitions = max_num_partitions;
  p.scale = scale;
  p.parts_per_chunk = 1;
  int n_chunks = 1;
  if (partitioned) {
    // Largest power-of-two chunk (in 512-token partitions) that still leaves >= 4
    // work items per resident CTA slot, so the last wave stays short.
    const int parts = (max_seq_len + kPartitionSize - 1) / kPartitionSize;
    const long slots = (long)num_sms * 2;
    const long per_part = (long)num_seqs * num_kv_heads * p.m_tiles;
    int ppc = 1;
    while (ppc < parts && per_part * ((parts + 2 * ppc - 1) / (2 * ppc)) >= 4 * slots) ppc *= 2;
    p.parts_per_chunk = ppc;
    n_chunks = (parts + ppc - 1) / ppc;
    if (n_chunks < 1) n_chunks = 1;
  }
  const bool fp8 = kv_dtype == B200PA_KV_FP8_E4M3;
#define FAST_CALL(ST, FP8V, DV) return launch_fast_t<ST, FP8V, DV>(p, num_seqs, n_chunks, partitioned, stream)
  if (scalar_type == B200PA_BF16) {
    if (fp8) { if (head_size == 128) FAST_CALL(B200PA_BF16, true, 128); else FAST_CALL(B200PA_BF16, true, 64); }
    else     { if (head_size == 128) FAST_CALL(B200PA_BF16, false, 128); else FAST_CALL(B200PA_BF16, false, 64); }
  } else {
    if (fp8) { if (head_size == 128) FAST_CALL(B200PA_F16, true, 128); else FAST_CALL(B200PA_F16, true, 64); }
    else     { if (head_size == 128) FAST_CALL(B200PA_F16, false, 128); else FAST_CALL(B200PA_F16, false, 64); }
  }
#undef FAST_CALL
}

}  // namespace b200pa
