// Causal variable-length attention of query CHUNKS over the paged KV cache (prefill, chunked
// prefill, multi-token decode) for sm_100a -- SURVEY 8f row 1.
//
// Replaces, for the paged backends of the reference, the second kernel they hand every batch with
// more than one query token per sequence to: flash_attn_varlen_func(out, query, key_cache,
// value_cache, cu_seqlens_q, cu_seqlens_kv, ..., causal, block_table, alibi_slopes)
// (vllm/v1/attention/backends/cpu_attn.py:584-612; the Triton context_attention_fwd behind
// PagedAttention.forward_prefix, vllm/attention/ops/paged_attn.py:200-242; the v1 flash-attention
// backend's call with block_table).  Semantics: sequence b has q_len = query_start_loc[b+1] -
// query_start_loc[b] new tokens whose K/V are ALREADY in the cache (the backends write the cache
// first); its query token i sits at position ctx + i, ctx = seq_lens[b] - q_len, and attends the
// cache positions [0, ctx + i] (bottom-right aligned causal mask).
//
//  * One CTA = one 64-row query tile of one (sequence, kv head): rows are (token, query head)
//    pairs -- `hpack` heads of the GQA group per token (1 for long chunks, the whole group for
//    short ones such as speculative decoding, so that the M tile is full and the context is
//    streamed once for the group).  4 warps x 16 rows.
//  * The KV side is the decode kernel's: the x-packed K block [D/x][BS][x] and the V block
//    [D][BS] of one (page, kv head) are contiguous runs, one 1-D TMA bulk copy each into a
//    CTA-wide ring of 64-token stages (full / empty mbarriers, warp 0 produces).  Block-table
//    entries are fetched 32 at a time, one window ahead.
//  * Math: mma.sync.m16n8k16, query rows on M: S = Q K^T with K fragments straight from the
//    x-packed layout (ldmatrix, no transposition needed: a 16-byte chunk IS a row of a B
//    fragment), O += P V with V fragments from the [D][BS] rows.  fp8 pages are converted in
//    registers (exact) and multiplied in f16, k_scale folded into the softmax scale, v_scale into
//    the output scale.  Online softmax in the log2 domain (FlashAttention-2 recurrence).
//  * Tiles of one (sequence, kv head) are adjacent in the grid, so they run concurrently and
//    share the context through the 126 MB L2: DRAM reads stay near one pass over the context.
#include "common.cuh"

namespace b200pa {

struct PrefillParams {
  void* out;                 // [T, H, D] scalar_t, token stride out_stride
  float* lse;                // optional [T, H] natural-log sum-exp
  const void* q;             // [T, H, D], token stride q_stride
  const void* k_cache;
  const void* v_cache;
  const int32_t* block_tables;
  const int32_t* query_start_loc;   // [num_seqs + 1]
  const int32_t* seq_lens;          // [num_seqs] context + new tokens
  const float* alibi_slopes;
  const float* k_scale;
  const float* v_scale;
  int64_t q_stride, out_stride, kv_block_stride, kv_head_stride;   // elements
  int num_seqs, num_heads, num_kv_heads, group;
  int max_num_blocks_per_seq;
  int hpack;                 // query heads of one kv head stacked per token in a tile (divides group)
  int tokens_per_tile;       // 64 / hpack
  float scale;
};

namespace {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr uint32_t kFull = 0xffffffffu;

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
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "PF_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra PF_DONE;\n"
      "bra PF_WAIT;\n"
      "PF_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void* src, uint32_t bytes,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(dst), "l"(src), "r"(bytes), "r"(bar)
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
// fp8 x4 (one 32-bit word) -> two f16x2 words: lo = bytes 0,1 ; hi = bytes 2,3 (exact)
template <bool E5>
__device__ __forceinline__ void fp8x4_to_f16x2x2(uint32_t w, uint32_t& lo, uint32_t& hi) {
  if constexpr (E5) {
    asm("prmt.b32 %0, %1, 0, 0x1404;" : "=r"(lo) : "r"(w));
    asm("prmt.b32 %0, %1, 0, 0x3424;" : "=r"(hi) : "r"(w));
  } else {
    asm("{\n"
        ".reg .b16 l, h;\n"
        "mov.b32 {l, h}, %2;\n"
        "cvt.rn.f16x2.e4m3x2 %0, l;\n"
        "cvt.rn.f16x2.e4m3x2 %1, h;\n"
        "}\n"
        : "=r"(lo), "=r"(hi)
        : "r"(w));
  }
}

constexpr int kTM = 64;        // query rows per CTA
constexpr int kTN = 64;        // KV tokens per pipeline stage
constexpr int kNWarps = 4;

// ST: scalar type of query / out.  KVT: B200PA_KV_* of the cache.  BSZ: tokens per cache block.
template <int ST, int KVT, int D, int BSZ, int STAGES>
__global__ void __launch_bounds__(kNWarps * 32, (D <= 128 ? 2 : 1))
paged_prefill_kernel(const PrefillParams p) {
  using T = typename Scalar<ST>::type;
  constexpr bool FP8 = KVT != B200PA_KV_AUTO;
  constexpr bool E5 = KVT == B200PA_KV_FP8_E5M2;
  constexpr bool MMA_BF16 = (!FP8) && (ST == B200PA_BF16);   // fp8 pages compute in f16
  constexpr int CES = FP8 ? 1 : 2;                  // cache element bytes
  constexpr int PPT = kTN / 16;                     // 16-token pages per stage
  constexpr int TPB = BSZ / 16;                     // pages per cache block
  constexpr int BPT = PPT / TPB;                    // cache blocks per stage
  constexpr int PAGE_BYTES = D * 16 * CES;
  constexpr int BLOCK_BYTES = TPB * PAGE_BYTES;     // one K (or V) block of one kv head
  constexpr int STAGE_BYTES = BPT * 2 * BLOCK_BYTES;
  constexpr int KCS = BSZ * 16;                     // x-packed K: bytes between 16-byte chunks
  constexpr int VRS = BSZ * CES;                    // x-packed V: bytes per head-element row
  constexpr int QROW = D * 2 + 16;                  // padded query row in shared memory
  constexpr int KSTEPS = D / 16;
  constexpr int DN = D / 8;                         // output n-tiles
  constexpr bool QREG = D <= 128;                   // query fragments live in registers
  static_assert(BSZ == 16 || BSZ == 32, "block size");
  static_assert(D % 32 == 0, "head size");

  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[2 * STAGES];   // full[STAGES], empty[STAGES]
  uint8_t* const qs = smem + STAGES * STAGE_BYTES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int mat = lane >> 3, mi = lane & 7;

  const int b = blockIdx.z;
  const int q_begin = p.query_start_loc[b];
  const int q_len = p.query_start_loc[b + 1] - q_begin;
  const int seq_len = p.seq_lens[b];
  const int tile = gridDim.x - 1 - blockIdx.x;      // long tiles first
  const int t0 = tile * p.tokens_per_tile;
  const int ctx = seq_len - q_len;                  // tokens before this chunk
  if (t0 >= q_len) return;
  const int hp = p.hpack;
  const int head0 = blockIdx.y * hp;
  if (ctx < 0) {
    // inconsistent metadata (fewer tokens in the sequence than in its chunk): nothing to attend
    // to is defined.  Leave a defined result instead of stale memory: zero rows, lse = -inf.
    using T0 = typename Scalar<ST>::type;
    T0* const o = reinterpret_cast<T0*>(p.out);
    const int nt = min(p.tokens_per_tile, q_len - t0);
    for (int i = threadIdx.x; i < nt * hp * D; i += blockDim.x) {
      const int tok = i / (hp * D), r = i % (hp * D);
      o[(int64_t)(q_begin + t0 + tok) * p.out_stride + (int64_t)head0 * D + r] =
          Scalar<ST>::from_float(0.f);
    }
    if (p.lse != nullptr)
      for (int i = threadIdx.x; i < nt * hp; i += blockDim.x)
        p.lse[(int64_t)(q_begin + t0 + i / hp) * p.num_heads + head0 + i % hp] = -INFINITY;
    return;
  }
  const int kvh = head0 / p.group;
  const int ntok = min(p.tokens_per_tile, q_len - t0);
  const int rows_valid = ntok * hp;
  const int kv_end = ctx + t0 + ntok;               // positions [0, kv_end) are visible to the tile
  const int n_tiles = (kv_end + kTN - 1) / kTN;
  const int total_pages = (kv_end + 15) / 16;

  const uint32_t ring_u32 = smem_u32(smem);
  const uint32_t qs_u32 = smem_u32(qs);
  const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[STAGES]);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, kNWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }

  // ---------------------------------------------------------------- query tile -> shared memory
  {
    constexpr int GPR = D / 8;   // 16-byte granules per row
    for (int idx = threadIdx.x; idx < kTM * GPR; idx += kNWarps * 32) {
      const int row = idx / GPR, c = idx - row * GPR;
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (row < rows_valid) {
        const int tok = row / hp, hd = head0 + row - tok * hp;
        v = *reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(p.q) +
                                            (int64_t)(q_begin + t0 + tok) * p.q_stride +
                                            (int64_t)hd * D + c * 8);
        if constexpr (FP8 && ST == B200PA_BF16) {
          uint32_t* w = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const __nv_bfloat162 x = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
            w[i] = pack2<false>(__bfloat162float(x.x), __bfloat162float(x.y));
          }
        }
      }
      *reinterpret_cast<uint4*>(qs + row * QROW + c * 16) = v;
    }
  }
  __syncthreads();   // barriers initialised, query tile in place

  // ---------------------------------------------------------------- producer (warp 0)
  const int64_t blk_stride_bytes = p.kv_block_stride * CES;
  const uint8_t* const kc =
      reinterpret_cast<const uint8_t*>(p.k_cache) + (int64_t)kvh * p.kv_head_stride * CES;
  const uint8_t* const vc =
      reinterpret_cast<const uint8_t*>(p.v_cache) + (int64_t)kvh * p.kv_head_stride * CES;
  const int32_t* const bt_row = p.block_tables + (int64_t)b * p.max_num_blocks_per_seq;
  int bt_win = -1, bt_cur = 0, bt_nxt = 0;   // 32 block-table entries of window bt_win, and the next
  auto bt_load = [&](int win) {
    const int idx = (win << 5) + lane;
    return idx < p.max_num_blocks_per_seq ? bt_row[idx] : 0;
  };
  auto produce = [&](int j) {   // every lane of warp 0
    const int s = j % STAGES;
    const int page0 = j * PPT;
    const int npg = min(PPT, total_pages - page0);
    const int blk0 = page0 / TPB, nblk = (npg + TPB - 1) / TPB;
    const int win = blk0 >> 5;   // BPT divides 32: a stage's blocks sit in one window
    if (win != bt_win) {
      bt_cur = (bt_win >= 0 && win == bt_win + 1) ? bt_nxt : bt_load(win);
      bt_win = win;
      bt_nxt = bt_load(win + 1);   // in flight until the next window is entered
    }
    int phys[BPT];
#pragma unroll
    for (int i = 0; i < BPT; ++i) phys[i] = __shfl_sync(kFull, bt_cur, (blk0 + i) & 31);
    if (lane == 0) {
      if (j >= STAGES) mbar_wait(empty0 + 8 * s, ((j / STAGES) - 1) & 1);
      const uint32_t bar = full0 + 8 * s;
      const uint32_t dst = ring_u32 + s * STAGE_BYTES;
      mbar_expect_tx(bar, nblk * 2 * BLOCK_BYTES);
#pragma unroll
      for (int i = 0; i < BPT; ++i) {
        if (i < nblk) {
          tma_load_1d(dst + i * 2 * BLOCK_BYTES, kc + (int64_t)phys[i] * blk_stride_bytes,
                      BLOCK_BYTES, bar);
          tma_load_1d(dst + i * 2 * BLOCK_BYTES + BLOCK_BYTES,
                      vc + (int64_t)phys[i] * blk_stride_bytes, BLOCK_BYTES, bar);
        }
      }
    }
    __syncwarp();
  };
  if (warp == 0) {
#pragma unroll 1
    for (int j = 0; j < STAGES - 1 && j < n_tiles; ++j) produce(j);
  }

  // ---------------------------------------------------------------- per-lane constants
  // accumulator rows of this lane: rowA = 16 warp + g, rowB = rowA + 8; row -> (token, head)
  const int rowA = 16 * warp + g, rowB = rowA + 8;
  const bool validA = rowA < rows_valid, validB = rowB < rows_valid;
  const int tokA = rowA / hp, tokB = rowB / hp;
  const int hdA = head0 + rowA - tokA * hp, hdB = head0 + rowB - tokB * hp;
  // largest visible cache position (rows past the tile see everything: never stored)
  const int limA = validA ? ctx + t0 + tokA : 0x7fffffff;
  const int limB = validB ? ctx + t0 + tokB : 0x7fffffff;
  const bool has_alibi = p.alibi_slopes != nullptr;
  float slopeA = 0.f, slopeB = 0.f;
  if (has_alibi) {
    if (validA) slopeA = p.alibi_slopes[hdA] * kLog2e;
    if (validB) slopeB = p.alibi_slopes[hdB] * kLog2e;
  }
  float sl2 = p.scale * kLog2e, out_scale = 1.f;
  if constexpr (FP8) {
    sl2 *= *p.k_scale;
    out_scale = *p.v_scale;
  }

  // query A-fragments of k-step ks.  16-bit caches: plain m16k16 fragments.  fp8 caches: the
  // k-slots (2t, 2t+1 | 2t+8, 2t+9) of a k-step hold head elements 16ks + (4t, 4t+1 | 4t+2, 4t+3),
  // the 4 consecutive bytes a lane gets from an fp8 K row via ldmatrix.
  const uint32_t q_lane = qs_u32 + (16 * warp + (mat & 1) * 8 + mi) * QROW + (mat >> 1) * 16;
  auto load_qfrag = [&](int ks, uint32_t (&a)[4]) {
    if constexpr (FP8) {
      const uint2 lo = *reinterpret_cast<const uint2*>(qs + rowA * QROW + (16 * ks + 4 * t) * 2);
      const uint2 hi = *reinterpret_cast<const uint2*>(qs + rowB * QROW + (16 * ks + 4 * t) * 2);
      a[0] = lo.x; a[2] = lo.y; a[1] = hi.x; a[3] = hi.y;
    } else {
      ldmatrix_x4(a, q_lane + ks * 32);
    }
  };
  uint32_t qf[QREG ? KSTEPS : 1][4];
  if constexpr (QREG) {
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) load_qfrag(ks, qf[ks]);
  }

  // ldmatrix lane offsets inside a page
  uint32_t k_lane_off, v_lane_off;
  int colA;   // fp8: token (within the page) of this lane's column pair in the page's first n-tile
  if constexpr (FP8) {
    // K x4 = {tokens A, chunk c | tokens B, chunk c | tokens A, chunk c+1 | tokens B, chunk c+1},
    // A: row i -> token 4(i/2) + i%2 + (i >= 4 ? 2 : 0), B: that ^ 2 (8 distinct bank groups each)
    const int tkA = 4 * (mi >> 1) + (mi & 1) + ((mi >= 4) ? 2 : 0);
    const int tk = (mat & 1) ? (tkA ^ 2) : tkA;
    k_lane_off = (mat >> 1) * KCS + tk * 16;
    // V x4 = output n-tiles dn .. dn+3: 8 head-element rows each, the tile's 16 tokens per row
    v_lane_off = (mat * 8 + mi) * VRS;
    colA = 4 * t + ((t >= 2) ? 2 : 0);
  } else {
    // K x4 = {chunk 2ks, tok 0-7 | chunk 2ks+1, tok 0-7 | chunk 2ks, tok 8-15 | chunk 2ks+1, tok 8-15}
    k_lane_off = (mat & 1) * KCS + ((mat >> 1) * 8 + mi) * 16;
    // V x4 = {dn, tok 0-7 | dn, tok 8-15 | dn+1, tok 0-7 | dn+1, tok 8-15}
    v_lane_off = ((mat >> 1) * 8 + mi) * VRS + (mat & 1) * 16;
    colA = 2 * t;
  }
  auto k_tile = [](int pg) { return (pg / TPB) * 2 * BLOCK_BYTES + (pg % TPB) * 256; };
  auto v_tile = [](int pg) { return (pg / TPB) * 2 * BLOCK_BYTES + BLOCK_BYTES + (pg % TPB) * 16 * CES; };
  // token (within its page) of column pair e of n-tile h (0 / 1) of a page, for this lane
  auto col_token = [&](int h) {
    if constexpr (FP8) return h ? (colA ^ 2) : colA;
    else return colA + 8 * h;
  };

  float o[DN][4];
#pragma unroll
  for (int dn = 0; dn < DN; ++dn) o[dn][0] = o[dn][1] = o[dn][2] = o[dn][3] = 0.f;
  float mA = -INFINITY, mB = -INFINITY, lA = 0.f, lB = 0.f;

  // ---------------------------------------------------------------- main loop over KV stages
#pragma unroll 1
  for (int j = 0; j < n_tiles; ++j) {
    if (warp == 0 && j + STAGES - 1 < n_tiles) produce(j + STAGES - 1);
    const int s = j % STAGES;
    mbar_wait(full0 + 8 * s, (j / STAGES) & 1);
    if (16 * warp >= rows_valid) {
      // all 16 rows of this warp are padding (a short chunk): keep in step, compute nothing
      __syncwarp();
      if (lane == 0) mbar_arrive(empty0 + 8 * s);
      continue;
    }
    const uint32_t sb = ring_u32 + s * STAGE_BYTES;
    const int npg = min(PPT, total_pages - j * PPT);   // pages of the stage that were loaded

    // ---- S = Q K^T : sc[2 pg + h] = 16 rows x 8 tokens (n-tile h of page pg)
    float sc[2 * PPT][4];
#pragma unroll
    for (int i = 0; i < 2 * PPT; ++i) sc[i][0] = sc[i][1] = sc[i][2] = sc[i][3] = 0.f;
    if constexpr (FP8) {
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ks += 2) {
        uint32_t a0[4], a1[4];
        if constexpr (QREG) {
#pragma unroll
          for (int i = 0; i < 4; ++i) { a0[i] = qf[ks][i]; a1[i] = qf[ks + 1][i]; }
        } else {
          load_qfrag(ks, a0);
          load_qfrag(ks + 1, a1);
        }
#pragma unroll
        for (int pg = 0; pg < PPT; ++pg) {
          if (pg < npg) {
            uint32_t w4[4], b0, b1;
            ldmatrix_x4(w4, sb + k_tile(pg) + ks * KCS + k_lane_off);
            fp8x4_to_f16x2x2<E5>(w4[0], b0, b1);
            mma16816<false>(sc[2 * pg], a0, b0, b1);
            fp8x4_to_f16x2x2<E5>(w4[1], b0, b1);
            mma16816<false>(sc[2 * pg + 1], a0, b0, b1);
            fp8x4_to_f16x2x2<E5>(w4[2], b0, b1);
            mma16816<false>(sc[2 * pg], a1, b0, b1);
            fp8x4_to_f16x2x2<E5>(w4[3], b0, b1);
            mma16816<false>(sc[2 * pg + 1], a1, b0, b1);
          }
        }
      }
    } else {
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        uint32_t a[4];
        if constexpr (QREG) {
#pragma unroll
          for (int i = 0; i < 4; ++i) a[i] = qf[ks][i];
        } else {
          load_qfrag(ks, a);
        }
#pragma unroll
        for (int pg = 0; pg < PPT; ++pg) {
          if (pg < npg) {
            uint32_t k4[4];
            ldmatrix_x4(k4, sb + k_tile(pg) + 2 * ks * KCS + k_lane_off);
            mma16816<MMA_BF16>(sc[2 * pg], a, k4[0], k4[1]);
            mma16816<MMA_BF16>(sc[2 * pg + 1], a, k4[2], k4[3]);
          }
        }
      }
    }

    // ---- scaled logits (log2 domain), ALiBi, causal mask
    const int kp0 = j * kTN;
    const bool need_mask = kp0 + kTN - 1 > ctx + t0;   // CTA-uniform: some row does not see it all
#pragma unroll
    for (int i = 0; i < 2 * PPT; ++i) {
      const int kp = kp0 + 16 * (i >> 1) + col_token(i & 1);   // position of element 0 (1: + 1)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int lim = (e & 2) ? limB : limA;
        const int pos = kp + (e & 1);
        float v = sc[i][e] * sl2;
        if (has_alibi) v += ((e & 2) ? slopeB : slopeA) * (float)(pos - lim);
        if (need_mask && pos > lim) v = -INFINITY;
        sc[i][e] = v;
      }
    }

    // ---- online softmax
    float mxA = -INFINITY, mxB = -INFINITY;
#pragma unroll
    for (int i = 0; i < 2 * PPT; ++i) {
      mxA = fmaxf(mxA, fmaxf(sc[i][0], sc[i][1]));
      mxB = fmaxf(mxB, fmaxf(sc[i][2], sc[i][3]));
    }
    mxA = fmaxf(mxA, __shfl_xor_sync(kFull, mxA, 1));
    mxA = fmaxf(mxA, __shfl_xor_sync(kFull, mxA, 2));
    mxB = fmaxf(mxB, __shfl_xor_sync(kFull, mxB, 1));
    mxB = fmaxf(mxB, __shfl_xor_sync(kFull, mxB, 2));
    const float nmA = fmaxf(mA, mxA), nmB = fmaxf(mB, mxB);
    // (a valid row sees position 0, so its maximum is finite from the first stage on)
    const float alA = ex2(mA - nmA), alB = ex2(mB - nmB);
    mA = nmA;
    mB = nmB;
    float sumA = 0.f, sumB = 0.f;
#pragma unroll
    for (int i = 0; i < 2 * PPT; ++i) {
      sc[i][0] = ex2(sc[i][0] - nmA);
      sc[i][1] = ex2(sc[i][1] - nmA);
      sc[i][2] = ex2(sc[i][2] - nmB);
      sc[i][3] = ex2(sc[i][3] - nmB);
      sumA += sc[i][0] + sc[i][1];
      sumB += sc[i][2] + sc[i][3];
    }
    lA = lA * alA + sumA;
    lB = lB * alB + sumB;
#pragma unroll
    for (int dn = 0; dn < DN; ++dn) {
      o[dn][0] *= alA; o[dn][1] *= alA;
      o[dn][2] *= alB; o[dn][3] *= alB;
    }

    // ---- O += P V, one k-step per page
    // the sequence's last page may hold anything (NaN bit patterns included) past seq_len: those
    // V columns are zeroed as the reference's decode kernel does (attention_kernels.cuh:415-425)
    const bool tail = (j == n_tiles - 1) && (total_pages * 16 > seq_len);
#pragma unroll
    for (int pg = 0; pg < PPT; ++pg) {
      if (pg < npg) {
        uint32_t pa[4];
        {
          const uint32_t x0 = pack2<MMA_BF16>(sc[2 * pg][0], sc[2 * pg][1]);
          const uint32_t x1 = pack2<MMA_BF16>(sc[2 * pg][2], sc[2 * pg][3]);
          const uint32_t y0 = pack2<MMA_BF16>(sc[2 * pg + 1][0], sc[2 * pg + 1][1]);
          const uint32_t y1 = pack2<MMA_BF16>(sc[2 * pg + 1][2], sc[2 * pg + 1][3]);
          // fp8: k-slots (2t, 2t+1) pair with V bytes 0,1 = tokens 4t, 4t+1, which lanes t >= 2
          // hold in the page's SECOND n-tile (see k_lane_off)
          const bool sw = FP8 && (t >= 2);
          pa[0] = sw ? y0 : x0;
          pa[1] = sw ? y1 : x1;
          pa[2] = sw ? x0 : y0;
          pa[3] = sw ? x1 : y1;
        }
        uint32_t m0 = 0xffffffffu, m1 = 0xffffffffu;   // masks of b0 / b1 (token pairs)
        if (tail && pg == npg - 1) {
          const int nv = seq_len - (total_pages - 1) * 16;        // valid tokens of the page
          const int c0 = FP8 ? 4 * t : 2 * t;                      // tokens of b0: c0, c0 + 1
          const int c1 = FP8 ? 4 * t + 2 : 2 * t + 8;              // tokens of b1
          m0 = (c0 < nv ? 0x0000ffffu : 0u) | (c0 + 1 < nv ? 0xffff0000u : 0u);
          m1 = (c1 < nv ? 0x0000ffffu : 0u) | (c1 + 1 < nv ? 0xffff0000u : 0u);
        }
        const uint32_t vb = sb + v_tile(pg) + v_lane_off;
        if constexpr (FP8) {
#pragma unroll
          for (int dn = 0; dn < DN; dn += 4) {
            uint32_t w4[4];
            ldmatrix_x4(w4, vb + dn * 8 * VRS);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint32_t b0, b1;
              fp8x4_to_f16x2x2<E5>(w4[i], b0, b1);
              mma16816<false>(o[dn + i], pa, b0 & m0, b1 & m1);
            }
          }
        } else {
#pragma unroll
          for (int dn = 0; dn < DN; dn += 2) {
            uint32_t v4[4];
            ldmatrix_x4(v4, vb + dn * 8 * VRS);
            mma16816<MMA_BF16>(o[dn], pa, v4[0] & m0, v4[1] & m1);
            mma16816<MMA_BF16>(o[dn + 1], pa, v4[2] & m0, v4[3] & m1);
          }
        }
      }
    }

    // ---- this warp is done with the stage
    __syncwarp();
    if (lane == 0) mbar_arrive(empty0 + 8 * s);
  }

  // ---------------------------------------------------------------- epilogue
  lA += __shfl_xor_sync(kFull, lA, 1);
  lA += __shfl_xor_sync(kFull, lA, 2);
  lB += __shfl_xor_sync(kFull, lB, 1);
  lB += __shfl_xor_sync(kFull, lB, 2);
  const float invA = lA > 0.f ? out_scale / lA : 0.f;
  const float invB = lB > 0.f ? out_scale / lB : 0.f;
  T* const outp = reinterpret_cast<T*>(p.out);
  constexpr bool OUT_BF16 = ST == B200PA_BF16;
  if (validA) {
    uint32_t* dst = reinterpret_cast<uint32_t*>(
        outp + (int64_t)(q_begin + t0 + tokA) * p.out_stride + (int64_t)hdA * D + 2 * t);
#pragma unroll
    for (int dn = 0; dn < DN; ++dn) dst[4 * dn] = pack2<OUT_BF16>(o[dn][0] * invA, o[dn][1] * invA);
  }
  if (validB) {
    uint32_t* dst = reinterpret_cast<uint32_t*>(
        outp + (int64_t)(q_begin + t0 + tokB) * p.out_stride + (int64_t)hdB * D + 2 * t);
#pragma unroll
    for (int dn = 0; dn < DN; ++dn) dst[4 * dn] = pack2<OUT_BF16>(o[dn][2] * invB, o[dn][3] * invB);
  }
  if (p.lse != nullptr && t == 0) {
    if (validA)
      p.lse[(int64_t)(q_begin + t0 + tokA) * p.num_heads + hdA] = (mA + __log2f(lA)) * kLn2;
    if (validB)
      p.lse[(int64_t)(q_begin + t0 + tokB) * p.num_heads + hdB] = (mB + __log2f(lB)) * kLn2;
  }
}

template <int ST, int KVT, int D, int BSZ>
int launch_prefill_t(const PrefillParams& p, int max_query_len, cudaStream_t stream) {
  constexpr bool FP8 = KVT != B200PA_KV_AUTO;
  constexpr int CES = FP8 ? 1 : 2;
  // ring depth: 64-token stages of 2 * 64 * D * CES bytes; two CTAs per SM up to head size 128
  constexpr int STAGES = FP8 ? 3 : (D <= 64 ? 3 : 2);
  constexpr int SMEM = STAGES * 2 * kTN * D * CES + kTM * (D * 2 + 16);
  auto kern = paged_prefill_kernel<ST, KVT, D, BSZ, STAGES>;
  static bool attr_set[64] = {false};   // the shared-memory opt-in is per device
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
    return set_error(B200PA_ERR_CUDA, "cudaGetDevice failed or device ordinal out of range");
  if (!attr_set[dev]) {
    const cudaError_t e =
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess)
      return set_error(B200PA_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    attr_set[dev] = true;
  }
  const int tiles = (max_query_len + p.tokens_per_tile - 1) / p.tokens_per_tile;
  const dim3 grid(tiles, p.num_heads / p.hpack, p.num_seqs);
  kern<<<grid, kNWarps * 32, SMEM, stream>>>(p);
  count_launch();
  return check_launch("prefill_attention");
}

}  // namespace

// The instantiations are spread over translation units by head size (this file compiled with
// -DPREFILL_PART=0..3, see build.py): 0 entry point + 128, 1: 32 / 64, 2: 96 / 192, 3: 256.
#ifndef PREFILL_PART
#define PREFILL_PART 0
#endif

#define PREFILL_BY_TYPES(DV)                                                                       \
  do {                                                                                             \
    if (scalar_type == B200PA_BF16) {                                                              \
      if (block_size == 16) {                                                                      \
        if (kv_dtype == B200PA_KV_AUTO) return launch_prefill_t<B200PA_BF16, B200PA_KV_AUTO, DV, 16>(p, max_query_len, stream); \
        if (kv_dtype == B200PA_KV_FP8_E4M3) return launch_prefill_t<B200PA_BF16, B200PA_KV_FP8_E4M3, DV, 16>(p, max_query_len, stream); \
        return launch_prefill_t<B200PA_BF16, B200PA_KV_FP8_E5M2, DV, 16>(p, max_query_len, stream); \
      }                                                                                            \
      if (kv_dtype == B200PA_KV_AUTO) return launch_prefill_t<B200PA_BF16, B200PA_KV_AUTO, DV, 32>(p, max_query_len, stream); \
      if (kv_dtype == B200PA_KV_FP8_E4M3) return launch_prefill_t<B200PA_BF16, B200PA_KV_FP8_E4M3, DV, 32>(p, max_query_len, stream); \
      return launch_prefill_t<B200PA_BF16, B200PA_KV_FP8_E5M2, DV, 32>(p, max_query_len, stream);  \
    }                                                                                              \
    if (block_size == 16) {                                                                        \
      if (kv_dtype == B200PA_KV_AUTO) return launch_prefill_t<B200PA_F16, B200PA_KV_AUTO, DV, 16>(p, max_query_len, stream); \
      if (kv_dtype == B200PA_KV_FP8_E4M3) return launch_prefill_t<B200PA_F16, B200PA_KV_FP8_E4M3, DV, 16>(p, max_query_len, stream); \
      return launch_prefill_t<B200PA_F16, B200PA_KV_FP8_E5M2, DV, 16>(p, max_query_len, stream);   \
    }                                                                                              \
    if (kv_dtype == B200PA_KV_AUTO) return launch_prefill_t<B200PA_F16, B200PA_KV_AUTO, DV, 32>(p, max_query_len, stream); \
    if (kv_dtype == B200PA_KV_FP8_E4M3) return launch_prefill_t<B200PA_F16, B200PA_KV_FP8_E4M3, DV, 32>(p, max_query_len, stream); \
    return launch_prefill_t<B200PA_F16, B200PA_KV_FP8_E5M2, DV, 32>(p, max_query_len, stream);     \
  } while (0)

int launch_prefill_small(const PrefillParams& p, int head_size, int block_size, int scalar_type,
                         int kv_dtype, int max_query_len, cudaStream_t stream);
int launch_prefill_mid(const PrefillParams& p, int head_size, int block_size, int scalar_type,
                       int kv_dtype, int max_query_len, cudaStream_t stream);
int launch_prefill_256(const PrefillParams& p, int head_size, int block_size, int scalar_type,
                       int kv_dtype, int max_query_len, cudaStream_t stream);

#if PREFILL_PART == 1
int launch_prefill_small(const PrefillParams& p, int head_size, int block_size, int scalar_type,
                         int kv_dtype, int max_query_len, cudaStream_t stream) {
  if (head_size == 32) PREFILL_BY_TYPES(32);
  PREFILL_BY_TYPES(64);
}
#elif PREFILL_PART == 2
int launch_prefill_mid(const PrefillParams& p, int head_size, int block_size, int scalar_type,
                       int kv_dtype, int max_query_len, cudaStream_t stream) {
  if (head_size == 96) PREFILL_BY_TYPES(96);
  PREFILL_BY_TYPES(192);
}
#elif PREFILL_PART == 3
int launch_prefill_256(const PrefillParams& p, int head_size, int block_size, int scalar_type,
                       int kv_dtype, int max_query_len, cudaStream_t stream) {
  PREFILL_BY_TYPES(256);
}
#else
bool prefill_supported(int head_size, int block_size, int scalar_type, int kv_dtype) {
  if (head_size != 32 && head_size != 64 && head_size != 96 && head_size != 128 &&
      head_size != 192 && head_size != 256) return false;
  if (block_size != 16 && block_size != 32) return false;
  if (scalar_type != B200PA_BF16 && scalar_type != B200PA_F16) return false;
  return kv_dtype == B200PA_KV_AUTO || kv_dtype == B200PA_KV_FP8_E4M3 ||
         kv_dtype == B200PA_KV_FP8_E5M2;
}

int launch_prefill(const PrefillParams& p, int head_size, int block_size, int scalar_type,
                   int kv_dtype, int max_query_len, cudaStream_t stream) {
  if (head_size == 128) PREFILL_BY_TYPES(128);
  if (head_size == 32 || head_size == 64)
    return launch_prefill_small(p, head_size, block_size, scalar_type, kv_dtype, max_query_len, stream);
  if (head_size == 96 || head_size == 192)
    return launch_prefill_mid(p, head_size, block_size, scalar_type, kv_dtype, max_query_len, stream);
  return launch_prefill_256(p, head_size, block_size, scalar_type, kv_dtype, max_query_len, stream);
}
#endif
#undef PREFILL_BY_TYPES

}  // namespace b200pa

#if PREFILL_PART == 0
using namespace b200pa;

extern "C" int b200pa_prefill_attention(
    void* out, float* lse, const void* query, const void* key_cache, const void* value_cache,
    int num_seqs, int num_heads, int num_kv_heads, int head_size, int block_size, float scale,
    const int32_t* block_tables, int max_num_blocks_per_seq, const int32_t* query_start_loc,
    const int32_t* seq_lens, int max_query_len, const float* alibi_slopes, int64_t q_stride,
    int64_t out_stride, int64_t kv_block_stride, int64_t kv_head_stride, int scalar_type,
    int kv_dtype, const float* k_scale, const float* v_scale, int heads_per_tile, void* stream) {
  if (num_seqs < 0 || num_heads <= 0 || num_kv_heads <= 0 || num_heads % num_kv_heads != 0)
    return set_error(B200PA_ERR_INVALID, "num_heads %d must be a positive multiple of num_kv_heads %d",
                     num_heads, num_kv_heads);
  if (!out || !query || !key_cache || !value_cache || !block_tables || !query_start_loc || !seq_lens)
    return set_error(B200PA_ERR_INVALID, "null tensor pointer");
  if (!prefill_supported(head_size, block_size, scalar_type, kv_dtype))
    return set_error(B200PA_ERR_UNSUPPORTED, "prefill_attention: head size %d / block size %d / "
                     "scalar type %d / kv dtype %d is outside the kernel's domain (head sizes 32, "
                     "64, 96, 128, 192, 256; block sizes 16, 32; f16 / bf16 query)", head_size,
                     block_size, scalar_type, kv_dtype);
  if (kv_dtype != B200PA_KV_AUTO && (!k_scale || !v_scale))
    return set_error(B200PA_ERR_INVALID, "fp8 kv cache needs k_scale and v_scale");
  const int ces = kv_dtype == B200PA_KV_AUTO ? 2 : 1;
  if ((reinterpret_cast<uintptr_t>(key_cache) | reinterpret_cast<uintptr_t>(value_cache)) & 15u ||
      (kv_block_stride * ces) % 16 != 0 || (kv_head_stride * ces) % 16 != 0)
    return set_error(B200PA_ERR_INVALID, "kv cache must be 16-byte aligned (base and strides)");
  if (kv_head_stride < (int64_t)head_size * block_size)
    return set_error(B200PA_ERR_INVALID, "kv_head_stride is smaller than one (block, head)");
  if ((reinterpret_cast<uintptr_t>(query) & 15u) || q_stride % 8 != 0)
    return set_error(B200PA_ERR_INVALID, "query rows must be 16-byte aligned");
  if ((reinterpret_cast<uintptr_t>(out) & 3u) || out_stride % 2 != 0)
    return set_error(B200PA_ERR_INVALID, "out rows must be 4-byte aligned");
  if (num_seqs == 0 || max_query_len <= 0) return B200PA_OK;
  if (num_seqs > 65535)
    return set_error(B200PA_ERR_UNSUPPORTED, "prefill_attention: more than 65535 sequences");
  const int group = num_heads / num_kv_heads;
  // heads of the GQA group stacked per token in a 64-row tile: the whole group for short chunks
  // (one pass over the context for all of them), one head per tile for long ones
  int hpack = heads_per_tile;
  if (hpack <= 0) hpack = (max_query_len < kTM && group <= kTM) ? group : 1;
  if (hpack > kTM || group % hpack != 0)
    return set_error(B200PA_ERR_INVALID, "heads_per_tile %d must divide the GQA group %d and be "
                     "at most %d", hpack, group, kTM);
  PrefillParams p;
  p.out = out;
  p.lse = lse;
  p.q = query;
  p.k_cache = key_cache;
  p.v_cache = value_cache;
  p.block_tables = block_tables;
  p.query_start_loc = query_start_loc;
  p.seq_lens = seq_lens;
  p.alibi_slopes = alibi_slopes;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.q_stride = q_stride;
  p.out_stride = out_stride;
  p.kv_block_stride = kv_block_stride;
  p.kv_head_stride = kv_head_stride;
  p.num_seqs = num_seqs;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.group = group;
  p.max_num_blocks_per_seq = max_num_blocks_per_seq;
  p.hpack = hpack;
  p.tokens_per_tile = kTM / hpack;
  p.scale = scale;
  return launch_prefill(p, head_size, block_size, scalar_type, kv_dtype, max_query_len,
                        reinterpret_cast<cudaStream_t>(stream));
}
#endif
