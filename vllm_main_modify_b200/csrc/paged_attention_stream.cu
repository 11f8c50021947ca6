// Persistent "stream-K" paged-decode attention for sm_100a (serves v1 AND v2).
//
// Replaces the per-sequence decode kernels of the reference
// (csrc/attention/attention_kernels.cuh:85-491 through paged_attention_v1_kernel :494-522 and
// paged_attention_v2_kernel :525-557 + paged_attention_v2_reduce_kernel :560-664, launched from
// paged_attention_v1.cu:46-125 / paged_attention_v2.cu:47-126) for head_size 32/64/96/128/192/256
// and 80 / 112 (x-packed layout),
// block_size 16/32, bf16/fp16 query and a 16-bit or fp8-e4m3 KV cache -- x-packed layout, or
// (FLASH, block_size 16, the same head sizes) the token-major layout of reshape_and_cache_flash --
// with the new token's cache write optionally fused in (decode_step).
//
//  * WORK DECOMPOSITION.  The unit of work is one pipeline STAGE: PPS consecutive pages of one
//    (sequence, kv head) -- 8 KiB of KV.  All stages of the call are flattened in (sequence,
//    kv head, page) order and cut into gridDim.x * NW equal contiguous spans, one per WARP.
//    The grid exactly fills the machine (SMs x resident CTAs) and every warp streams its span
//    from start to end, so all warps finish together whatever the mix of sequence lengths.
//    Ragged batches need no host-side knowledge of seq_lens: each warp derives its span from
//    seq_lens on the device (CUDA-graph replays with new lengths just work).
//  * Each warp is an independent worker: its own TMA ring, mbarriers, query fragments and
//    accumulators.  There is no CTA-wide barrier in the main loop.  When a span crosses into
//    another (sequence, kv head) the warp finishes the piece it was on and carries on; the TMA
//    producer side runs STAGES ahead across such boundaries, so the pipe never drains.
//    A piece that covers its whole (sequence, kv head) is normalised and written straight to
//    `out`.  A warp has at most two pieces that do not (its first and its last): their
//    unnormalised fp32 accumulators + (max, sum) go to workspace slots 2w / 2w+1.  At the end of
//    the kernel -- behind its first CTA-wide barrier -- each CTA merges the pieces of its own
//    warps: items that begin and end inside the CTA are finished there, and what is cut by a CTA
//    boundary (at most two runs per CTA) goes to the CTA's slots 2c / 2c+1.  The CTAs that share
//    such an item take a ticket from the item's counter in the workspace; the LAST one to arrive
//    merges all the item's CTA slots (all its warps, fixed piece order: deterministic whoever is
//    last), writes `out` and resets the counter -- the job of the reference's
//    paged_attention_v2_reduce_kernel (attention_kernels.cuh:560-664) without a second launch.
//    (Round 1 used a separate merge kernel: 7-11 us per step, profiles/r01q_*.)
//  * LOADS.  For one (page, kv head) the x-packed K sub-block [D/x][16][x] and the V sub-block
//    [D][16] are each one contiguous run (2 KiB fp8 / 4 KiB 16-bit at D=128): one 1-D TMA bulk
//    copy each, completing on the stage's mbarrier.  Block-table entries are fetched 32 at a
//    time per warp.
//  * MATH.  Tokens sit on the M axis of mma.sync.m16n8k16 and the (<= 8) query heads of the GQA
//    group on the N axis: S^T = K Q^T and O^T = V^T P^T.  One page is exactly one M tile, so
//    the query group is not padded to 16 rows; P^T reaches the B-operand layout with one
//    movmatrix per 8x8 block.  fp8 pages are converted in registers (cvt.rn.f16x2.e4m3x2,
//    exact) and multiplied in f16, k_scale folded into the softmax scale and v_scale into the
//    output scale.  Online softmax in the log2 domain with a lazily moved reference: the
//    steady state has no shuffles, only a warp vote.
#include <cuda.h>   // CUtensorMap (types only; the encoder is fetched through the runtime)

#include "common.cuh"

// Development builds only (tools/ablate.sh): B200PA_ABLATE removes one ingredient of the main loop
// so that its share of the step's time and POWER can be measured (results are then wrong):
// 1 no fp8 -> f16 conversion, 2 no MMA, 4 no ldmatrix, 8 no exp2.  0 in every product build.
#ifndef B200PA_ABLATE
#define B200PA_ABLATE 0
#endif

namespace b200pa {

namespace {

constexpr int kBS = 16;                      // tokens per page (block_size)
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr uint32_t kFull = 0xffffffffu;

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
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void* src, uint32_t bytes,
                                            uint32_t bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar), "l"(policy)
      : "memory");
}
// 16-byte asynchronous copy global -> shared (LDGSTS), L1 bypassed, L2 evict-first
__device__ __forceinline__ void cp_async_16(uint32_t dst, const void* src, uint64_t policy) {
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(dst), "l"(src),
               "l"(policy)
               : "memory");
}
// this thread's earlier cp.async copies arrive on the mbarrier when they have landed (the
// arrival counts against the barrier's expected count: .noinc)
__device__ __forceinline__ void cp_async_mbar_arrive(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
#if B200PA_ABLATE & 4
  r[0] = addr; r[1] = addr + 1; r[2] = addr + 2; r[3] = addr + 3;
#else
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
#endif
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
template <uint32_t SEL>
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "n"(SEL));
  return d;
}
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}
template <bool BF16>
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                         uint32_t b1) {
#if B200PA_ABLATE & 2
  asm volatile("" ::"r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  d[0] += 1.f;
  return;
#endif
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
#if B200PA_ABLATE & 8
  return x;
#endif
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// weight 2^(m - mx) of a partial whose reference is m: a piece that saw no key at all for a head
// (block-sparse attention: every block of the piece masked) has m = -inf and weight 0, also when
// the whole run is like that (mx = -inf)
__device__ __forceinline__ float piece_weight(float m, float mx) {
  return m == -INFINITY ? 0.f : ex2(m - mx);
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
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
// fp8 x4 (one 32-bit word) -> two f16x2 words: lo = bytes 0,1 ; hi = bytes 2,3.  Exact for both
// formats; an e5m2 code is the high byte of the f16 with the same value (two byte permutes).
template <bool E5>
__device__ __forceinline__ void fp8x4_to_f16x2x2(uint32_t w, uint32_t& lo, uint32_t& hi) {
  if constexpr (E5) {
    asm("prmt.b32 %0, %1, 0, 0x1404;" : "=r"(lo) : "r"(w));
    asm("prmt.b32 %0, %1, 0, 0x3424;" : "=r"(hi) : "r"(w));
    return;
  }
#if B200PA_ABLATE & 1
  lo = w; hi = w;
  return;
#endif
  asm("{\n"
      ".reg .b16 l, h;\n"
      "mov.b32 {l, h}, %2;\n"
      "cvt.rn.f16x2.e4m3x2 %0, l;\n"
      "cvt.rn.f16x2.e4m3x2 %1, h;\n"
      "}\n"
      : "=r"(lo), "=r"(hi)
      : "r"(w));
}


}  // namespace

// (named, not file-local: the launchers of the four translation-unit parts below share it)
struct StreamParams {
  void* out;           // [S,H,D] scalar_t
  float* lse;          // optional: natural-log sum-exp of the scaled logits per (seq, head)
  int64_t lse_seq_stride, lse_head_stride;
  float* ws;           // workspace slots: 2 per worker + 2 per CTA, slot_floats(D) floats each
  int* counters;       // workspace header: one ticket counter per CTA, zero between launches
  // development aid (b200pa_testing_set_trace, nullptr otherwise): 32 %globaltimer stamps per CTA
  // -- 0 entry, 1 dependency wait over, 2 stage count known, 3 span located, 4 first copies
  // issued, 5 query loaded, 8+w first stage landed (warp w), 16+w main loop over (warp w),
  // 24 every warp over, 25 in-CTA merge over, 26 ticket taken, 27 exit
  unsigned long long* trace;
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const float* alibi_slopes;
  const float* k_scale;
  const float* v_scale;
  // fused cache write (decode_step), new_key == nullptr: plain attention.  The warp that
  // consumes the last stage of a (sequence, kv head) patches the new token into its shared-memory
  // copy of the page (every head tile) and stores it to the global cache (head tile 0 only).
  const void* new_key;
  const void* new_value;
  const int64_t* slot_mapping;
  int64_t new_key_stride, new_value_stride;           // elements per token
  int64_t q_stride, kv_block_stride, kv_head_stride;  // elements
  int64_t kv_token_stride;   // flash layout only: elements between two tokens of a block
  int num_seqs, num_heads, num_kv_heads, group;       // group = num_heads / num_kv_heads
  int h_tiles;                                         // ceil(group / 8)
  int max_num_blocks_per_seq;
  // block-sparse attention (attention_kernels.cuh:207-251 of the reference); vert_stride <= 1: off
  int bs_vert_stride, bs_local_blocks, bs_block_size, bs_head_sliding_step, bs_tp_rank;
  int num_workers;                                     // gridDim.x * NW of the main kernel
  int nw;                                              // NW: warps per CTA of the main kernel
  float scale;
};

namespace {

// workspace slot: O^T fragments [D/16][32 lanes][4] fp32 (unnormalised, v_scale applied),
// then m[8] (log2 domain), l[8]
__host__ __device__ constexpr int slot_floats(int d) { return 8 * d + 16; }
// workspace header: the ticket counters of the cross-CTA merge, one int per CTA.  ZERO before the
// first launch on a workspace and left zero by every launch; at a fixed offset so that launches
// of different shapes can share one workspace.
constexpr int kCounterBytes = 4096;

// flash layout: bytes of a token row of one head in shared memory -- its data plus at least 16
// bytes, an ODD number of 16-byte units so that the 8 rows of an ldmatrix fall into 8 different
// bank groups
__host__ __device__ constexpr int flash_row_bytes(int row_data) {
  return ((row_data + 16) / 16) % 2 ? row_data + 16 : row_data + 32;
}

template <int PPS> __device__ __forceinline__ int stages_of(int len) {
  return len <= 0 ? 0 : (len + kBS * PPS - 1) / (kBS * PPS);
}

// Sum of stages over all sequences (one warp; every lane gets the result).
template <int PPS>
__device__ __forceinline__ long long total_stages(const int32_t* seq_lens, int num_seqs, int lane) {
  int tot = 0;
  for (int b = lane; b < num_seqs; b += 32) tot += stages_of<PPS>(seq_lens[b]);
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) tot += __shfl_xor_sync(kFull, tot, m);
  return tot;
}

// Position of flattened stage index s (order: sequence, virtual head, stage).
struct Loc { int b, vh, len, ns, st; };
template <int PPS>
__device__ __forceinline__ Loc locate_stage(const int32_t* seq_lens, int num_seqs, int vheads,
                                            long long s, int lane) {
  Loc loc{0, 0, 0, 1, 0};
  int base = 0;
  for (int c = 0; c < num_seqs; c += 32) {
    const int len = (c + lane < num_seqs) ? seq_lens[c + lane] : 0;
    const int n = stages_of<PPS>(len);
    int incl = n;
#pragma unroll
    for (int m = 1; m < 32; m <<= 1) {
      const int v = __shfl_up_sync(kFull, incl, m);
      if (lane >= m) incl += v;
    }
    const int chunk_total = __shfl_sync(kFull, incl, 31);
    if ((long long)(base + chunk_total) * vheads > s) {
      const unsigned mask = __ballot_sync(kFull, (long long)(base + incl) * vheads > s);
      const int l = __ffs(mask) - 1;
      const int excl = __shfl_sync(kFull, incl - n, l);
      loc.b = c + l;
      loc.len = __shfl_sync(kFull, len, l);
      loc.ns = __shfl_sync(kFull, n, l);
      const long long rem = s - (long long)(base + excl) * vheads;
      loc.vh = (int)(rem / loc.ns);
      loc.st = (int)(rem - (long long)loc.vh * loc.ns);
      break;
    }
    base += chunk_total;
  }
  return loc;
}

// Where a worker is in its span.  A "piece" is the run of consecutive stages of one
// (sequence, virtual kv head) inside the span.
struct Cursor {
  int b, vh;        // sequence, virtual kv head (kv head x query-head tile)
  int len, ns;      // seq_lens[b], its stages
  int st;           // next stage within the (sequence, head)
  int st_end;       // end of the current piece (exclusive)
  int left;         // stages of the span beyond the current piece
};
__device__ __forceinline__ void cursor_open(Cursor& c, int st0) {
  const int n = min(c.left, c.ns - st0);
  c.st = st0;
  c.st_end = st0 + n;
  c.left -= n;
}
// precondition: c.left > 0 (so another non-empty sequence / head exists)
template <int PPS>
__device__ __forceinline__ void cursor_next_piece(Cursor& c, const int32_t* seq_lens, int vheads) {
  if (++c.vh == vheads) {
    c.vh = 0;
    do {
      ++c.b;
      c.len = seq_lens[c.b];
      c.ns = stages_of<PPS>(c.len);
    } while (c.ns == 0);
  }
  cursor_open(c, 0);
}

// Final output of one (sequence, 8-head tile) from O^T fragments: scale, round to scalar_t,
// transpose each 8x8 (d, head) block with movmatrix so that every lane stores 32-bit pairs.
// o[mt][j]: j = 0,1 -> d = 16mt + r, heads 2q, 2q+1 ; j = 2,3 -> d = 16mt + r + 8.
// DPERM (fp8 pages of the flash layout): the m-tile's rows are permuted, o[mt][0,1] holding
// d = 16mt + 2r and o[mt][2,3] d = 16mt + 2r + 1 -- after the transposes lane (r, q) holds head r,
// d = 16mt + 4q .. 4q+3: one 8-byte store.
template <int ST, int D, bool DPERM>
__device__ __forceinline__ void store_out_tile(typename Scalar<ST>::type* out_tile,  // [8 heads][D]
                                               int mt, const float (&o)[4], float inv0, float inv1,
                                               int heads_valid, int r, int q4) {
  constexpr bool BF = (ST == B200PA_BF16);
  const uint32_t lo = movmatrix_trans(pack2<BF>(o[0] * inv0, o[1] * inv1));
  const uint32_t hi = movmatrix_trans(pack2<BF>(o[2] * inv0, o[3] * inv1));
  if constexpr (DPERM) {
    if (r < heads_valid)
      *reinterpret_cast<uint2*>(out_tile + (int64_t)r * D + 16 * mt + 4 * q4) =
          make_uint2(prmt<0x5410>(lo, hi), prmt<0x7632>(lo, hi));
  } else {
    uint32_t* dst = reinterpret_cast<uint32_t*>(out_tile + (int64_t)r * D + 2 * q4);
    if (r < heads_valid) {          // after the transpose lane (r, q) holds head r
      dst[8 * mt] = lo;             // d = 16mt + 2q, 2q+1
      dst[8 * mt + 4] = hi;         // d = 16mt + 8 + 2q, 2q+1
    }
  }
}
template <int ST, int D, bool DPERM>
__device__ __forceinline__ void store_out(typename Scalar<ST>::type* out_tile,
                                          const float (&o)[D / 16][4], float inv0, float inv1,
                                          int heads_valid, int r, int q4) {
#pragma unroll
  for (int mt = 0; mt < D / 16; ++mt)
    store_out_tile<ST, D, DPERM>(out_tile, mt, o[mt], inv0, inv1, heads_valid, r, q4);
}

// ST: B200PA_F16 | B200PA_BF16 (query/out type).  FP8: kv cache is e4m3 bytes.
// PPS: pages per pipeline stage.
// FLASH: the cache is token-major, [block][token][head][D] by strides (NHD or HND, what
// reshape_and_cache_flash writes): every token row of D elements is one bulk copy into a shared
// row padded by 16 bytes (conflict-free ldmatrix), K fragments come from plain and V fragments
// from transposing ldmatrix.
// BSZ: tokens per cache block (16 or 32).  The kernel's unit stays the 16-token M tile ("page"
// below); with BSZ = 32 a cache block is two tiles that arrive in one bulk copy and are addressed
// inside the block by the layout's strides (K chunk stride BSZ*16 B, V row stride BSZ elements).
// E5: the fp8 codes are e5m2 (kv_cache_dtype "fp8_e5m2") instead of e4m3.
template <int ST, bool FP8, int D, int NW, int STAGES, int PPS, int MINB, bool FLASH, int BSZ,
          bool E5>
__global__ void __launch_bounds__(NW * 32, MINB)
paged_attention_stream_kernel(const StreamParams p) {
  static_assert(FP8 || !E5, "E5 selects the fp8 format");
  using T = typename Scalar<ST>::type;
  constexpr bool MMA_BF16 = (!FP8) && (ST == B200PA_BF16);  // fp8 pages compute in f16
  constexpr int CES = FP8 ? 1 : 2;                           // cache element bytes
  constexpr int PAGE_BYTES = D * kBS * CES;                  // one K (or V) page of one kv head
  constexpr int ROW_DATA = D * CES;                          // flash: one token row of one head
  constexpr int ROWB = flash_row_bytes(ROW_DATA);            // ... and its padded shared-memory row
  constexpr int PAGE_SMEM = FLASH ? kBS * ROWB : PAGE_BYTES; // shared-memory footprint of a page
  constexpr int STAGE_BYTES = PPS * 2 * PAGE_SMEM;
  constexpr bool DPERM = FLASH && FP8;                       // see store_out_tile
  constexpr int TPB = BSZ / kBS;                             // 16-token tiles per cache block
  constexpr int BPS = PPS / TPB;                             // cache blocks per stage
  constexpr int BLOCK_BYTES = TPB * PAGE_BYTES;              // one K (or V) block of one kv head
  constexpr int BLOCK_SMEM = TPB * PAGE_SMEM;
  constexpr int KCS = BSZ * 16;                              // x-packed K: bytes between 16-byte chunks
  constexpr int VRS = BSZ * CES;                             // x-packed V: bytes per head-element row
  static_assert(BSZ == 16 || BSZ == 32, "block size");
  static_assert(PPS % TPB == 0 && 32 % BPS == 0, "a stage holds whole cache blocks");
  // tile g of a stage: block g / TPB, half g % TPB (x-packed: inside the block's K / V run;
  // flash layout: every 16-token tile has its own K and V region of padded rows)
  auto k_tile = [](int g) {
    return FLASH ? g * 2 * PAGE_SMEM : (g / TPB) * 2 * BLOCK_SMEM + (g % TPB) * (kBS * 16);
  };
  auto v_tile = [](int g) {
    return FLASH ? g * 2 * PAGE_SMEM + PAGE_SMEM
                 : (g / TPB) * 2 * BLOCK_SMEM + BLOCK_SMEM + (g % TPB) * (kBS * CES);
  };
  constexpr int KSTEPS = D / 16;                             // QK k-steps
  constexpr int MT = D / 16;                                 // PV m-tiles (16 head elements each)
  constexpr int RING_BYTES = STAGES * STAGE_BYTES;
  // probabilities are exp2(t - msub) with t - msub <= CAP; msub trails the max by OFF
  constexpr float OFF = MMA_BF16 ? 0.f : 6.f;
  constexpr float CAP = MMA_BF16 ? 8.f : 14.f;

  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[NW * STAGES];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = lane >> 2, q4 = lane & 3;
  const int vheads = p.num_kv_heads * p.h_tiles;
#define TRACE(k) do { if (p.trace != nullptr && threadIdx.x == 0) \
    p.trace[(size_t)blockIdx.x * 32 + (k)] = globaltimer_ns(); } while (0)
#define TRACE_W(k) do { if (p.trace != nullptr && lane == 0) \
    p.trace[(size_t)blockIdx.x * 32 + (k) + warp] = globaltimer_ns(); } while (0)
  TRACE(0);

  // Launched with programmatic stream serialisation: this CTA may have been scheduled while the
  // previous kernel of the stream (typically reshape_and_cache writing this step's token) was
  // still running.  Nothing produced by earlier kernels -- seq_lens, q, the cache -- is read
  // before this wait; what overlaps is the launch itself.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  TRACE(1);

  // ---------------------------------------------------------------- my span
  const long long Ttot = total_stages<PPS>(p.seq_lens, p.num_seqs, lane) * vheads;
  TRACE(2);
  const long long W = min((long long)p.num_workers, Ttot);
  const long long w = (long long)blockIdx.x * NW + warp;
  // let the next kernel of the stream (the next layer's decode step) be scheduled as soon as SM
  // resources free up; it blocks in its own griddepcontrol.wait until this grid has completed
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // what this warp leaves for the in-CTA merge: bit 0 head piece (slot 2w, continues an earlier
  // worker's item), bit 1 the head piece reaches the item's end, bit 2 tail piece (slot 2w+1,
  // from the item's start, continued by a later worker)
  __shared__ int meta_s[NW][7];   // flags, head (b, vh), tail (b, vh), stages of the head / tail item
  __shared__ long long ibeg_s[NW][2];  // first flattened stage of the head / tail piece's item
  __shared__ float hdr_s[2 * NW][16];  // per piece (2i head, 2i+1 tail): m[8] (log2), l[8]
  // what the in-CTA merge leaves for the cross-CTA merge: [0] the run that continues an earlier
  // CTA's item (CTA slot 2c), [1] the run a later CTA continues (slot 2c+1)
  __shared__ int pend_s[2][6];    // valid, first CTA of the item, CTAs of the item, b, vh, last?
  int mflags = 0, mh_b = 0, mh_vh = 0, mt_b = 0, mt_vh = 0, mh_ns = 0, mt_ns = 0;
  long long mh_ib = 0, mt_ib = 0;
  if (threadIdx.x < 2) pend_s[threadIdx.x][0] = 0;
  if (w < W) {  // idle warps (fewer stages than warps) go straight to the barrier
  const long long u0 = (w * Ttot) / W, u1 = ((w + 1) * Ttot) / W;

  Cursor pc;  // producer cursor; the consumer gets a copy
  {
    const Loc loc = locate_stage<PPS>(p.seq_lens, p.num_seqs, vheads, u0, lane);
    pc.b = loc.b; pc.vh = loc.vh; pc.len = loc.len; pc.ns = loc.ns;
    pc.left = (int)(u1 - u0);
    cursor_open(pc, loc.st);
  }
  TRACE(3);
  Cursor cc = pc;
  int piece_begin = cc.st;
  int it = 0;                               // stages consumed

  // ---------------------------------------------------------------- ring + barriers
  uint8_t* ring = smem + (size_t)warp * RING_BYTES;
  const uint32_t ring_u32 = smem_u32(ring);
  const uint32_t bar0 = smem_u32(&bars[warp * STAGES]);
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(bar0 + 8 * s, FLASH ? 32 : 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();

  const int64_t blk_stride_bytes = p.kv_block_stride * CES;
  const int64_t head_stride_bytes = p.kv_head_stride * CES;
  const uint64_t policy = l2_evict_first_policy();

  // ---------------------------------------------------------------- producer
  int bt_reg = 0, bt_b = -1, bt_win = -1;  // 32 block-table entries of (bt_b, window bt_win)
  bool producing = true;
  int ip = 0;                               // stages issued
  auto produce = [&]() {
    if (pc.st == pc.st_end) {
      if (pc.left == 0) { producing = false; return; }
      cursor_next_piece<PPS>(pc, p.seq_lens, vheads);
    }
    const int page0 = pc.st * PPS;
    const int npg = min(PPS, (pc.len + kBS - 1) / kBS - page0);
    ++pc.st;
    const int blk0 = page0 / TPB;                 // first cache block of the stage
    const int nblk = (npg + TPB - 1) / TPB;       // blocks with at least one valid tile
    const int win = blk0 >> 5;
    if (pc.b != bt_b || win != bt_win) {
      const int idx = (win << 5) + lane;
      bt_reg = (idx < p.max_num_blocks_per_seq)
                   ? p.block_tables[(int64_t)pc.b * p.max_num_blocks_per_seq + idx] : 0;
      bt_b = pc.b;
      bt_win = win;
    }
    int phys[BPS];
#pragma unroll
    for (int g = 0; g < BPS; ++g) phys[g] = __shfl_sync(kFull, bt_reg, (blk0 + g) & 31);
    if constexpr (FLASH) {
      // Token rows of one head are D elements apart from each other by the token stride: no
      // single run to bulk-copy.  (One TMA bulk copy per 128/256-byte row reached 3.0 TB/s on
      // every shape, profiles/r01e_kbench_flash.jsonl.)  16-byte cp.async granules instead,
      // dealt to the lanes so that a warp instruction covers whole rows; every lane arrives
      // once per stage on the stage's mbarrier (expected count 32).
      constexpr int GR = ROW_DATA / 16;            // granules per token row
      constexpr int PER_LANE = (kBS * GR + 31) / 32;   // granules per lane per K (or V) page
      const int s = ip % STAGES;
      const uint32_t bar = bar0 + 8 * s;
      const int kvh = pc.vh / p.h_tiles;
      const int64_t tsb = p.kv_token_stride * CES;
      const uint8_t* kc = reinterpret_cast<const uint8_t*>(p.k_cache) + kvh * head_stride_bytes;
      const uint8_t* vc = reinterpret_cast<const uint8_t*>(p.v_cache) + kvh * head_stride_bytes;
      const uint32_t dst = ring_u32 + s * STAGE_BYTES;
#pragma unroll
      for (int g = 0; g < PPS; ++g) {
        if (g < npg) {
          // tile g: block g / TPB of the stage, tokens (g % TPB) * 16 .. + 15 of it
          const int64_t boff = (int64_t)phys[g / TPB] * blk_stride_bytes + (g % TPB) * kBS * tsb;
#pragma unroll
          for (int i = 0; i < PER_LANE; ++i) {
            const int idx = i * 32 + lane;
            if ((kBS * GR) % 32 == 0 || idx < kBS * GR) {
              const int row = idx / GR, col = idx % GR;
              const int64_t goff = boff + row * tsb + col * 16;
              const uint32_t soff = g * 2 * PAGE_SMEM + row * ROWB + col * 16;
              cp_async_16(dst + soff, kc + goff, policy);
              cp_async_16(dst + soff + PAGE_SMEM, vc + goff, policy);
            }
          }
        }
      }
      cp_async_mbar_arrive(bar);
    } else if (lane == 0) {
      const int s = ip % STAGES;
      const uint32_t bar = bar0 + 8 * s;
      const uint32_t dst = ring_u32 + s * STAGE_BYTES;
      const int kvh = pc.vh / p.h_tiles;
      const uint8_t* kc = reinterpret_cast<const uint8_t*>(p.k_cache) + kvh * head_stride_bytes;
      const uint8_t* vc = reinterpret_cast<const uint8_t*>(p.v_cache) + kvh * head_stride_bytes;
      mbar_expect_tx(bar, nblk * 2 * BLOCK_BYTES);
#pragma unroll
      for (int g = 0; g < BPS; ++g) {
        if (g < nblk) {
          tma_load_1d(dst + g * 2 * BLOCK_BYTES, kc + (int64_t)phys[g] * blk_stride_bytes,
                      BLOCK_BYTES, bar, policy);
          tma_load_1d(dst + g * 2 * BLOCK_BYTES + BLOCK_BYTES,
                      vc + (int64_t)phys[g] * blk_stride_bytes, BLOCK_BYTES, bar, policy);
        }
      }
    }
    ++ip;
  };
#pragma unroll 1
  for (int i = 0; i < STAGES && producing; ++i) produce();
  TRACE(4);

  // ---------------------------------------------------------------- consumer state
  uint32_t qb[KSTEPS][2];       // Q^T B-fragments: head r, k-step ks
  float o[MT][4];               // O^T: (d = 16mt + r | + 8, heads 2q4, 2q4+1)
  float msub[2], lsum[2];       // per head 2q4 + h: subtracted reference (log2), partial sum
  float slope_l2[2];
  int head0 = 0, heads_valid = 0;
  const bool has_alibi = p.alibi_slopes != nullptr;
  // block-sparse: per query head the offset of its vertical stripes, per piece the query's block
  const bool block_sparse = p.bs_vert_stride > 1;
  int bs_off[2] = {0, 0}, bs_qblk = 0;
  float sl2 = p.scale * kLog2e;
  float out_scale = 1.f;
  float ks_raw = 1.f;           // k_scale itself (the fused cache write quantises with it)
  if constexpr (FP8) {
    ks_raw = *p.k_scale;
    sl2 *= ks_raw;
    out_scale = *p.v_scale;
  }
  // fused cache write: this lane's pairs of head elements (2(lane + 32j), +1) of the new token's
  // K and V (raw scalar_t pairs) and its slot, fetched when a piece that reaches the end of its
  // sequence begins
  constexpr int NP = (D / 2 + 31) / 32;
  uint32_t nt_k[NP], nt_v[NP];
  int64_t nt_slot = -1;

  // ldmatrix lane addressing
  const int mat = lane >> 3, mi = lane & 7;
  uint32_t k_lane_off, v_lane_off;
  int tokA, tokB;  // tokens (within the page) of accumulator rows r and r + 8
  if constexpr (FLASH) {
    // K x4 (plain) = {tok 0-7, 16 B | tok 8-15, same 16 B | tok 0-7, next 16 B | tok 8-15, next}
    // V x4 (transposing): 16-bit {tok 0-7, d 0-7 | tok 0-7, d 8-15 | tok 8-15, d 0-7 | tok 8-15,
    // d 8-15}; fp8 {tok 0-7, d 0-15 | tok 8-15, d 0-15 | tok 0-7, d 16-31 | tok 8-15, d 16-31}
    k_lane_off = ((mat & 1) * 8 + mi) * ROWB + (mat >> 1) * 16;
    v_lane_off = FP8 ? k_lane_off : ((mat >> 1) * 8 + mi) * ROWB + (mat & 1) * 16;
    tokA = r;
    tokB = r + 8;
  } else if constexpr (FP8) {
    // K x4 = {chunk c, slots 0-7 | chunk c, slots 8-15 | chunk c+1, slots 0-7 | chunk c+1, 8-15}
    // (chunk = 16 head elements).  Slot i < 8 is token sig0(i) = 4(i/2) + (i%2) + (i >= 4 ? 2 : 0)
    // and slot 8 + i is token sig0(i) ^ 2: both halves touch 8 distinct 16-byte bank groups, and
    // after the movmatrix transpose quad-lane q holds P for tokens 4q..4q+3 -- the 4 bytes it
    // gets from a V row (b0/b1 swapped for q >= 2).
    const int sig0 = 4 * (mi >> 1) + (mi & 1) + ((mi >= 4) ? 2 : 0);
    const int tok = (mat & 1) ? (sig0 ^ 2) : sig0;
    k_lane_off = (mat >> 1) * KCS + tok * 16;
    v_lane_off = lane * VRS;  // 32 consecutive V rows (head elements) per x4
    tokA = 4 * (r >> 1) + (r & 1) + ((r >= 4) ? 2 : 0);
    tokB = tokA ^ 2;
  } else {
    // K x4 = {chunk 2ks, tok 0-7 | chunk 2ks, tok 8-15 | chunk 2ks+1, tok 0-7 | chunk 2ks+1, tok 8-15}
    // (chunk = 8 head elements)
    k_lane_off = (mat >> 1) * KCS + ((mat & 1) * 8 + mi) * 16;
    // V x4 = {d 0-7, tok 0-7 | d 8-15, tok 0-7 | d 0-7, tok 8-15 | d 8-15, tok 8-15}
    v_lane_off = ((mat & 1) * 8 + mi) * VRS + (mat >> 1) * 16;
    tokA = r;
    tokB = r + 8;
  }

  auto begin_piece = [&]() {
    piece_begin = cc.st;
    const int kvh = cc.vh / p.h_tiles;
    if (p.new_key != nullptr) {
      nt_slot = -1;
      if (cc.st_end == cc.ns) {
        nt_slot = p.slot_mapping[cc.b];
        const uint32_t* kn = reinterpret_cast<const uint32_t*>(
            reinterpret_cast<const T*>(p.new_key) + (int64_t)cc.b * p.new_key_stride + (int64_t)kvh * D);
        const uint32_t* vn = reinterpret_cast<const uint32_t*>(
            reinterpret_cast<const T*>(p.new_value) + (int64_t)cc.b * p.new_value_stride + (int64_t)kvh * D);
#pragma unroll
        for (int j = 0; j < NP; ++j) {
          const int pi = lane + 32 * j;
          nt_k[j] = nt_v[j] = 0u;
          if (pi < D / 2) {
            nt_k[j] = kn[pi];
            nt_v[j] = vn[pi];
          }
        }
      }
    }
    const int tile = cc.vh - kvh * p.h_tiles;
    head0 = kvh * p.group + tile * 8;
    heads_valid = min(8, p.group - tile * 8);
    const T* qrow = reinterpret_cast<const T*>(p.q) + (int64_t)cc.b * p.q_stride +
                    (int64_t)(head0 + r) * D;
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) {
      uint32_t v0 = 0, v1 = 0;
      if (r < heads_valid) {
        if constexpr (FP8) {
          // k-slots (2q,2q+1 | 2q+8,2q+9) of k-step ks hold head elements 16ks + (4q,4q+1 |
          // 4q+2,4q+3): the 4 consecutive bytes a lane gets from an fp8 K row via ldmatrix
          const uint2 raw = *reinterpret_cast<const uint2*>(qrow + 16 * ks + 4 * q4);
          if constexpr (ST == B200PA_BF16) {
            const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
            const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
            v0 = pack2<false>(__bfloat162float(a.x), __bfloat162float(a.y));
            v1 = pack2<false>(__bfloat162float(b.x), __bfloat162float(b.y));
          } else {
            v0 = raw.x;
            v1 = raw.y;
          }
        } else {
          v0 = *reinterpret_cast<const uint32_t*>(qrow + 16 * ks + 2 * q4);
          v1 = *reinterpret_cast<const uint32_t*>(qrow + 16 * ks + 8 + 2 * q4);
        }
      }
      qb[ks][0] = v0;
      qb[ks][1] = v1;
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      slope_l2[h] = 0.f;
      if (has_alibi && 2 * q4 + h < heads_valid)
        slope_l2[h] = p.alibi_slopes[head0 + 2 * q4 + h] * kLog2e;
      msub[h] = -INFINITY;
      lsum[h] = 0.f;
    }
    if (block_sparse) {
      // attention_kernels.cuh:209-220: stripes slide with the query head (step >= 0) or the kv head
      bs_qblk = (cc.len - 1) / p.bs_block_size;
#pragma unroll
      for (int h = 0; h < 2; ++h)
        bs_off[h] = p.bs_head_sliding_step >= 0
            ? (p.bs_tp_rank * p.num_heads + head0 + 2 * q4 + h) * p.bs_head_sliding_step + 1
            : (p.bs_tp_rank * p.num_kv_heads + kvh) * (-p.bs_head_sliding_step) + 1;
    }
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) o[mt][0] = o[mt][1] = o[mt][2] = o[mt][3] = 0.f;
  };

  auto end_piece = [&]() {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      lsum[h] += __shfl_xor_sync(kFull, lsum[h], 4);
      lsum[h] += __shfl_xor_sync(kFull, lsum[h], 8);
      lsum[h] += __shfl_xor_sync(kFull, lsum[h], 16);
    }
    const bool from_start = piece_begin == 0, to_end = cc.st_end == cc.ns;
    if (from_start && to_end) {
      // the piece is the whole (sequence, head tile): final result
      const float inv0 = out_scale * __fdividef(1.f, lsum[0] + 1e-6f);
      const float inv1 = out_scale * __fdividef(1.f, lsum[1] + 1e-6f);
      store_out<ST, D, DPERM>(reinterpret_cast<T*>(p.out) + ((int64_t)cc.b * p.num_heads + head0) * D, o,
                       inv0, inv1, heads_valid, r, q4);
      if (p.lse != nullptr && r == 0) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
          if (2 * q4 + h < heads_valid)
            p.lse[cc.b * p.lse_seq_stride + (head0 + 2 * q4 + h) * p.lse_head_stride] =
                (msub[h] + __log2f(lsum[h])) * kLn2;
      }
    } else {
      // slot 2w: my first piece, continuing an earlier worker's; slot 2w+1: my last piece,
      // continued by a later worker
      // the item's first flattened stage: this piece ends at stage u0 + it of the span
      const long long ib = u0 + it - cc.st_end;
      if (from_start) { mflags |= 4; mt_b = cc.b; mt_vh = cc.vh; mt_ns = cc.ns; mt_ib = ib; }
      else { mflags |= to_end ? 3 : 1; mh_b = cc.b; mh_vh = cc.vh; mh_ns = cc.ns; mh_ib = ib; }
      float* slot = p.ws + (size_t)(2 * w + (from_start ? 1 : 0)) * slot_floats(D);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
        reinterpret_cast<float4*>(slot)[mt * 32 + lane] = make_float4(
            o[mt][0] * out_scale, o[mt][1] * out_scale, o[mt][2] * out_scale, o[mt][3] * out_scale);
      if (r == 0) {
        float* hdr = hdr_s[2 * warp + (from_start ? 1 : 0)];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          hdr[2 * q4 + h] = msub[h];
          hdr[8 + 2 * q4 + h] = lsum[h];
        }
      }
    }
  };

  begin_piece();
  TRACE(5);

  // ---------------------------------------------------------------- main loop
#pragma unroll 1
  for (;; ++it) {
    const int i = it;
    if (cc.st == cc.st_end) {
      end_piece();
      if (cc.left == 0) break;
      cursor_next_piece<PPS>(cc, p.seq_lens, vheads);
      begin_piece();
    }
    const int page0 = cc.st * PPS;
    const int npg = min(PPS, (cc.len + kBS - 1) / kBS - page0);
    ++cc.st;
    const int s = i % STAGES;
    const uint32_t sbase = ring_u32 + s * STAGE_BYTES;
    mbar_wait(bar0 + 8 * s, (i / STAGES) & 1);
    if (i == 0) TRACE_W(8);

    // ---- fused cache write: the stage that ends the sequence holds position len - 1, whose
    // K/V come from this step's key/value rows (in registers since begin_piece, quantised here
    // exactly as reshape_and_cache does), not from the bytes the TMA copy found in the cache.
    if (p.new_key != nullptr && cc.st == cc.ns && nt_slot >= 0) {
      constexpr int X = 16 / CES;       // elements per 16-byte chunk of the x-packed K layout
      const int t = cc.len - 1 - page0 * kBS;
      // x-packed: the token's cache block in the stage and its offset in it; flash: its 16-token
      // tile and its row in it
      uint8_t* pg = ring + s * STAGE_BYTES +
                    (FLASH ? (t >> 4) * (2 * PAGE_SMEM) : ((t >> 4) / TPB) * (2 * BLOCK_SMEM));
      const int tt = FLASH ? (t & (kBS - 1)) : (t & (BSZ - 1));
      const int kvh = cc.vh / p.h_tiles;
      const bool to_global = cc.vh == kvh * p.h_tiles;
      const int64_t goff = (nt_slot / BSZ) * blk_stride_bytes + kvh * head_stride_bytes;
      const int off = (int)(nt_slot % BSZ);
      uint8_t* kg = reinterpret_cast<uint8_t*>(const_cast<void*>(p.k_cache)) + goff;
      uint8_t* vg = reinterpret_cast<uint8_t*>(const_cast<void*>(p.v_cache)) + goff;
      // a pair of cache elements: 2 bytes (fp8) or 4 bytes (16-bit)
      auto put = [&](uint8_t* dst, uint32_t w) {
        if constexpr (FP8) *reinterpret_cast<uint16_t*>(dst) = (uint16_t)w;
        else *reinterpret_cast<uint32_t*>(dst) = w;
      };
#pragma unroll
      for (int j = 0; j < NP; ++j) {
        const int e0 = 2 * (lane + 32 * j);      // head elements e0, e0 + 1
        if (e0 < D) {
          uint32_t kw = nt_k[j], vw = nt_v[j];
          if constexpr (FP8) {
            const T* ke = reinterpret_cast<const T*>(&nt_k[j]);
            const T* ve = reinterpret_cast<const T*>(&nt_v[j]);
            kw = fp8_quant2<(E5 ? B200PA_KV_FP8_E5M2 : B200PA_KV_FP8_E4M3)>(Scalar<ST>::to_float(ke[0]),
                                                Scalar<ST>::to_float(ke[1]), ks_raw);
            vw = fp8_quant2<(E5 ? B200PA_KV_FP8_E5M2 : B200PA_KV_FP8_E4M3)>(Scalar<ST>::to_float(ve[0]),
                                                Scalar<ST>::to_float(ve[1]), out_scale);
          }
          if constexpr (FLASH) {
            // token-major rows: K and V of the token are contiguous runs
            const int64_t tsb = p.kv_token_stride * CES;
            put(pg + tt * ROWB + e0 * CES, kw);
            put(pg + PAGE_SMEM + tt * ROWB + e0 * CES, vw);
            if (to_global) {
              put(kg + off * tsb + e0 * CES, kw);
              put(vg + off * tsb + e0 * CES, vw);
            }
          } else {
            const int kchunk = (e0 / X) * KCS + (e0 % X) * CES;    // + token * 16
            put(pg + kchunk + tt * 16, kw);
            if (to_global) put(kg + kchunk + off * 16, kw);
#pragma unroll
            for (int i = 0; i < 2; ++i) {   // V [D][16]: one element per head-element row
              if constexpr (FP8) {
                const uint8_t b8 = (uint8_t)(vw >> (8 * i));
                pg[BLOCK_SMEM + (e0 + i) * BSZ + tt] = b8;
                if (to_global) vg[(e0 + i) * BSZ + off] = b8;
              } else {
                const uint16_t v16 = (uint16_t)(vw >> (16 * i));
                reinterpret_cast<uint16_t*>(pg + BLOCK_SMEM)[(e0 + i) * BSZ + tt] = v16;
                if (to_global) reinterpret_cast<uint16_t*>(vg)[(e0 + i) * BSZ + off] = v16;
              }
            }
          }
        }
      }
      // generic-proxy stores to a buffer the TMA engine refills later
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
    }

    // ---- S^T = K Q^T : per page one 16(token) x 8(head) tile, two k chains
    float sc[PPS][4];
#pragma unroll
    for (int g = 0; g < PPS; ++g) {
      sc[g][0] = sc[g][1] = sc[g][2] = sc[g][3] = 0.f;
      if (g < npg) {
        const uint32_t kbase = sbase + k_tile(g);
        float sb[4] = {0.f, 0.f, 0.f, 0.f};
        if constexpr (FP8) {
#pragma unroll
          for (int c = 0; c < KSTEPS; c += 2) {
            // (odd KSTEPS, D = 80 / 112: the second half of the last x4 reads the first bytes of
            // the V tile that follows the K tile in the stage -- valid memory, not used)
            uint32_t w4[4], a[4];
            ldmatrix_x4(w4, kbase + c * (FLASH ? 16 : KCS) + k_lane_off);
            fp8x4_to_f16x2x2<E5>(w4[0], a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(w4[1], a[1], a[3]);
            mma16816<false>(sc[g], a, qb[c][0], qb[c][1]);
            if (c + 1 < KSTEPS) {
              constexpr int KL = KSTEPS - 1;
              const int c1 = c + 1 < KL ? c + 1 : KL;
              fp8x4_to_f16x2x2<E5>(w4[2], a[0], a[2]);
              fp8x4_to_f16x2x2<E5>(w4[3], a[1], a[3]);
              mma16816<false>(sb, a, qb[c1][0], qb[c1][1]);
            }
          }
        } else {
#pragma unroll
          for (int ks = 0; ks < KSTEPS; ++ks) {
            uint32_t a[4];
            ldmatrix_x4(a, kbase + ks * (FLASH ? 32 : 2 * KCS) + k_lane_off);
            if (ks & 1) mma16816<MMA_BF16>(sb, a, qb[ks][0], qb[ks][1]);
            else mma16816<MMA_BF16>(sc[g], a, qb[ks][0], qb[ks][1]);
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) sc[g][j] += sb[j];
      }
    }

    // ---- online softmax.  d[g][j]: j = 0,1 token tokA heads 2q,2q+1 ; j = 2,3 token tokB
    const int tok0 = page0 * kBS;
    float d[PPS][4];
    auto logits = [&](float neg0, float neg1) {
#pragma unroll
      for (int g = 0; g < PPS; ++g) {
        const int nv = cc.len - (tok0 + g * kBS);  // valid tokens of the page (<= 0: absent)
        bool seen[2] = {true, true};
        if (block_sparse) {
          // attention_kernels.cuh:234-240: a cache block is attended by a head if it is one of
          // the last local blocks or on one of the head's vertical stripes.  (The reference skips
          // the other blocks; here they are streamed and masked -- same result.)
          const int kb = ((page0 + g) / TPB) * BSZ / p.bs_block_size;
          const bool local = kb > bs_qblk - p.bs_local_blocks;
          seen[0] = local || ((kb + bs_off[0]) % p.bs_vert_stride == 0);
          seen[1] = local || ((kb + bs_off[1]) % p.bs_vert_stride == 0);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int tk = (j & 2) ? tokB : tokA;
          float t = fmaf(sc[g][j], sl2, (j & 1) ? neg1 : neg0);
          if (has_alibi) t += slope_l2[j & 1] * (float)(tok0 + g * kBS + tk - cc.len + 1);
          d[g][j] = (g < npg && tk < nv && seen[j & 1]) ? t : -INFINITY;
        }
      }
    };
    logits(-msub[0], -msub[1]);
    bool over = false;
#pragma unroll
    for (int g = 0; g < PPS; ++g)
#pragma unroll
      for (int j = 0; j < 4; ++j) over |= d[g][j] > CAP;
    if (__any_sync(kFull, over)) {
      // the running max of some head outgrew its reference: exact maxima, move, rescale
      logits(0.f, 0.f);
      float alpha[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float mx = -INFINITY;
#pragma unroll
        for (int g = 0; g < PPS; ++g) mx = fmaxf(mx, fmaxf(d[g][h], d[g][2 + h]));
        mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, 4));
        mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, 8));
        mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, 16));
        alpha[h] = 1.f;
        if (mx > msub[h] + CAP) {
          const float nm = mx - OFF;
          alpha[h] = ex2(msub[h] - nm);  // ex2(-inf) = 0 on the first stage
          msub[h] = nm;
        }
        lsum[h] *= alpha[h];
      }
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        o[mt][0] *= alpha[0]; o[mt][1] *= alpha[1];
        o[mt][2] *= alpha[0]; o[mt][3] *= alpha[1];
      }
      logits(-msub[0], -msub[1]);
    }
    uint32_t pb[PPS][2];
#pragma unroll
    for (int g = 0; g < PPS; ++g) {
      float pr[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) pr[j] = ex2(d[g][j]);
      lsum[0] += pr[0] + pr[2];
      lsum[1] += pr[1] + pr[3];
      // 8x8 blocks (token slot x head) -> transpose -> B fragments (k = token, n = head)
      const uint32_t ta = movmatrix_trans(pack2<MMA_BF16>(pr[0], pr[1]));
      const uint32_t tb = movmatrix_trans(pack2<MMA_BF16>(pr[2], pr[3]));
      if constexpr (FP8 && !FLASH) {
        pb[g][0] = (q4 >= 2) ? tb : ta;
        pb[g][1] = (q4 >= 2) ? ta : tb;
      } else {
        pb[g][0] = ta;
        pb[g][1] = tb;
      }
    }

    // ---- O^T += V^T P^T
    auto pv = [&](int g, auto masked) {
      constexpr bool MASKED = decltype(masked)::value;
      const uint32_t vbase = sbase + v_tile(g);
      // tail page: V columns past the context may hold NaN bit patterns; the reference
      // zeroes them (attention_kernels.cuh:415-425).  m01 masks a[0], a[1]; m23 a[2], a[3].
      uint32_t m01 = 0xffffffffu, m23 = 0xffffffffu;
      if constexpr (MASKED) {
        const int nv = cc.len - (tok0 + g * kBS);
        const int t01 = (FP8 && !FLASH) ? 4 * q4 : 2 * q4;
        const int t23 = (FP8 && !FLASH) ? 4 * q4 + 2 : 2 * q4 + 8;
        m01 = (t01 < nv ? 0x0000ffffu : 0u) | (t01 + 1 < nv ? 0xffff0000u : 0u);
        m23 = (t23 < nv ? 0x0000ffffu : 0u) | (t23 + 1 < nv ? 0xffff0000u : 0u);
      }
      if constexpr (FP8 && FLASH) {
        // transposed 8x8 blocks of byte PAIRS: w4[0] = tokens (2q, 2q+1) x d (2r, 2r+1) of the
        // first 16 d, w4[1] the same for tokens + 8, w4[2], w4[3] for the next 16 d.  Rows of
        // the m-tile are assigned d = 2r (a[0], a[2]) and d = 2r + 1 (a[1], a[3]): DPERM.
#pragma unroll
        for (int gg = 0; gg < (D + 31) / 32; ++gg) {
          // (D = 80 / 112: the last group has 16 head elements; the second half of its x4 reads
          // the row's 16 bytes of padding -- valid memory, not used)
          uint32_t w4[4], a[4];
          ldmatrix_x4_trans(w4, vbase + gg * 32 + v_lane_off);
          fp8x4_to_f16x2x2<E5>(prmt<0x6420>(w4[0], w4[1]), a[0], a[2]);
          fp8x4_to_f16x2x2<E5>(prmt<0x7531>(w4[0], w4[1]), a[1], a[3]);
          if constexpr (MASKED) { a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23; }
          mma16816<false>(o[2 * gg], a, pb[g][0], pb[g][1]);
          if (32 * gg + 16 < D) {
            fp8x4_to_f16x2x2<E5>(prmt<0x6420>(w4[2], w4[3]), a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(prmt<0x7531>(w4[2], w4[3]), a[1], a[3]);
            if constexpr (MASKED) { a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23; }
            mma16816<false>(o[2 * gg + 1 < MT ? 2 * gg + 1 : MT - 1], a, pb[g][0], pb[g][1]);
          }
        }
      } else if constexpr (FP8) {
#pragma unroll
        for (int gg = 0; gg < (D + 31) / 32; ++gg) {
          uint32_t w4[4], a[4];
          // D = 80 / 112: the last group has 16 rows; lanes 16-31 re-read rows 0-15 of it (their
          // matrices are not used) instead of running past the V tile
          const bool half = 32 * gg + 16 >= D;
          ldmatrix_x4(w4, vbase + gg * (32 * VRS) + (half ? (lane & 15) * VRS : v_lane_off));
          fp8x4_to_f16x2x2<E5>(w4[0], a[0], a[2]);
          fp8x4_to_f16x2x2<E5>(w4[1], a[1], a[3]);
          if constexpr (MASKED) { a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23; }
          mma16816<false>(o[2 * gg], a, pb[g][0], pb[g][1]);
          if (!half) {
            constexpr int ML = MT - 1;
            const int m1 = 2 * gg + 1 < ML ? 2 * gg + 1 : ML;
            fp8x4_to_f16x2x2<E5>(w4[2], a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(w4[3], a[1], a[3]);
            if constexpr (MASKED) { a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23; }
            mma16816<false>(o[m1], a, pb[g][0], pb[g][1]);
          }
        }
      } else {
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          uint32_t a[4];
          if constexpr (FLASH) ldmatrix_x4_trans(a, vbase + mt * 32 + v_lane_off);
          else ldmatrix_x4(a, vbase + mt * (16 * VRS) + v_lane_off);
          if constexpr (MASKED) { a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23; }
          mma16816<MMA_BF16>(o[mt], a, pb[g][0], pb[g][1]);
        }
      }
    };
    if (tok0 + npg * kBS <= cc.len) {   // every page of the stage is full
#pragma unroll
      for (int g = 0; g < PPS; ++g)
        if (g < npg) pv(g, std::false_type{});
    } else {
#pragma unroll
      for (int g = 0; g < PPS; ++g)
        if (g < npg) pv(g, std::true_type{});
    }

    // ---- refill this stage (all lanes are past their last read of it)
    __syncwarp();
    if (producing) produce();
  }
  TRACE_W(16);
  }  // w < W

  // ---------------------------------------------------------------- in-CTA merge
  // Pieces in worker order: index 2i (head of warp i), 2i+1 (tail of warp i).  A run = the
  // consecutive pieces of one (sequence, head tile).  Every warp walks the same runs (metadata in
  // shared memory, so all branches are warp-uniform); run j is summed by warp j % NW, pieces in
  // index order (deterministic), two pieces' loads in flight at a time.  The (m, l) headers come
  // from shared memory, so the only global round trip is the one for the O fragments.  A run that
  // starts at the item's start and reaches its end is final; one that continues an earlier CTA's
  // item goes to CTA slot 2c, one that is continued by a later CTA to 2c+1 (same convention as
  // the warp slots).
  if (lane == 0) {
    meta_s[warp][0] = mflags;
    meta_s[warp][1] = mh_b; meta_s[warp][2] = mh_vh;
    meta_s[warp][3] = mt_b; meta_s[warp][4] = mt_vh;
    meta_s[warp][5] = mh_ns; meta_s[warp][6] = mt_ns;
    ibeg_s[warp][0] = mh_ib; ibeg_s[warp][1] = mt_ib;
  }
  // empty sequences: the reference's v1 kernel yields zeros (no keys, zero accumulators)
  for (long long b = (long long)blockIdx.x * NW + warp; b < p.num_seqs; b += (long long)gridDim.x * NW) {
    if (p.seq_lens[b] <= 0) {
      T* row = reinterpret_cast<T*>(p.out) + b * p.num_heads * D;
      for (int i = lane; i < p.num_heads * D; i += 32) row[i] = Scalar<ST>::from_float(0.f);
      if (p.lse != nullptr)
        for (int i = lane; i < p.num_heads; i += 32)
          p.lse[b * p.lse_seq_stride + i * p.lse_head_stride] = -INFINITY;
    }
  }
  __syncthreads();   // also orders the warps' global slot stores before the loads below
  TRACE(24);
  {
    constexpr int SF = slot_floats(D);
    const float* const my_slots = p.ws + (size_t)2 * ((size_t)blockIdx.x * NW) * SF;
    float* const cta_slots = p.ws + ((size_t)2 * p.num_workers + (size_t)2 * blockIdx.x) * SF;
    uint32_t pmask = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) {
      const int f = meta_s[i][0];
      pmask |= (uint32_t)(f & 1) << (2 * i);
      pmask |= (uint32_t)((f >> 2) & 1) << (2 * i + 1);
    }
    int run = 0;
    auto emit = [&](int ps, int pe, bool closed) {
      if ((run++ % NW) != warp) return;
      const bool from_start = ps & 1;
      const uint32_t mk = pmask & (0xffffffffu >> (31 - pe)) & (0xffffffffu << ps);
      float mx0 = -INFINITY, mx1 = -INFINITY;
      for (uint32_t t = mk; t; t &= t - 1) {
        const float* h = hdr_s[__ffs(t) - 1];
        mx0 = fmaxf(mx0, h[2 * q4]);
        mx1 = fmaxf(mx1, h[2 * q4 + 1]);
      }
      float4 acc[MT];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) acc[mt] = make_float4(0.f, 0.f, 0.f, 0.f);
      float l0 = 0.f, l1 = 0.f;
      for (uint32_t t = mk; t;) {
        const int pa = __ffs(t) - 1;
        t &= t - 1;
        const bool two = t != 0;
        const int pb = two ? __ffs(t) - 1 : pa;
        t &= t - 1;   // (0 & anything stays 0)
        const float4* sa = reinterpret_cast<const float4*>(my_slots + (size_t)pa * SF);
        const float4* sb = reinterpret_cast<const float4*>(my_slots + (size_t)pb * SF);
        float4 va[MT], vb[MT];
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) va[mt] = sa[mt * 32 + lane];
        if (two) {
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) vb[mt] = sb[mt * 32 + lane];
        }
        const float* ha = hdr_s[pa];
        const float fa0 = piece_weight(ha[2 * q4], mx0), fa1 = piece_weight(ha[2 * q4 + 1], mx1);
        l0 += ha[8 + 2 * q4] * fa0;
        l1 += ha[8 + 2 * q4 + 1] * fa1;
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          acc[mt].x += va[mt].x * fa0; acc[mt].y += va[mt].y * fa1;
          acc[mt].z += va[mt].z * fa0; acc[mt].w += va[mt].w * fa1;
        }
        if (two) {
          const float* hb = hdr_s[pb];
          const float fb0 = piece_weight(hb[2 * q4], mx0), fb1 = piece_weight(hb[2 * q4 + 1], mx1);
          l0 += hb[8 + 2 * q4] * fb0;
          l1 += hb[8 + 2 * q4 + 1] * fb1;
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            acc[mt].x += vb[mt].x * fb0; acc[mt].y += vb[mt].y * fb1;
            acc[mt].z += vb[mt].z * fb0; acc[mt].w += vb[mt].w * fb1;
          }
        }
      }
      if (closed && from_start) {
        const int b = meta_s[ps >> 1][3], vh = meta_s[ps >> 1][4];
        const int kvh = vh / p.h_tiles, tile = vh - kvh * p.h_tiles;
        const int hd0 = kvh * p.group + tile * 8, hv = min(8, p.group - tile * 8);
        const float inv0 = __fdividef(1.f, l0 + 1e-6f), inv1 = __fdividef(1.f, l1 + 1e-6f);
        T* dst = reinterpret_cast<T*>(p.out) + ((int64_t)b * p.num_heads + hd0) * D;
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          const float o4[4] = {acc[mt].x, acc[mt].y, acc[mt].z, acc[mt].w};
          store_out_tile<ST, D, DPERM>(dst, mt, o4, inv0, inv1, hv, r, q4);
        }
        if (p.lse != nullptr && r == 0) {
          if (2 * q4 < hv)
            p.lse[b * p.lse_seq_stride + (hd0 + 2 * q4) * p.lse_head_stride] =
                (mx0 + __log2f(l0)) * kLn2;
          if (2 * q4 + 1 < hv)
            p.lse[b * p.lse_seq_stride + (hd0 + 2 * q4 + 1) * p.lse_head_stride] =
                (mx1 + __log2f(l1)) * kLn2;
        }
      } else {
        float* dst = cta_slots + (from_start ? SF : 0);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) reinterpret_cast<float4*>(dst)[mt * 32 + lane] = acc[mt];
        if (r == 0) {
          *reinterpret_cast<float2*>(dst + 8 * D + 2 * q4) = make_float2(mx0, mx1);
          *reinterpret_cast<float2*>(dst + 8 * D + 8 + 2 * q4) = make_float2(l0, l1);
        }
        __threadfence();   // the slot is visible device-wide before this CTA takes its ticket
        if (lane == 0) {
          // the CTAs that hold a piece of this item: those of the workers of its first and last
          // stage (worker of stage s = the largest x with (x * Ttot) / W <= s)
          const int i0 = ps >> 1, hk = ps & 1;
          const long long ib = ibeg_s[i0][hk], ie = ib + meta_s[i0][5 + hk];
          const long long c_first = (((ib + 1) * W - 1) / Ttot) / NW;
          const long long c_last = ((ie * W - 1) / Ttot) / NW;
          int* pd = pend_s[from_start ? 1 : 0];
          pd[1] = (int)c_first;
          pd[2] = (int)(c_last - c_first) + 1;
          pd[3] = meta_s[i0][hk ? 3 : 1];
          pd[4] = meta_s[i0][hk ? 4 : 2];
          pd[0] = 1;
        }
      }
    };
    int ps = -1;
#pragma unroll 1
    for (int i = 0; i < NW; ++i) {
      const int f = meta_s[i][0];
      if (f & 1) {
        if (ps < 0) ps = 2 * i;
        if (f & 2) { emit(ps, 2 * i, true); ps = -1; }
      }
      if (f & 4) ps = 2 * i + 1;
    }
    if (ps >= 0) emit(ps, 2 * NW - 1, false);
  }

  // ---------------------------------------------------------------- cross-CTA merge
  // An item that CTA boundaries cut has one piece per CTA it touches: slot 2c+1 of its first CTA,
  // slot 2c of every later one.  Each of those CTAs takes a ticket from the counter of the item's
  // first CTA; the last to arrive sums the pieces with the reference's weights l_j * 2^(m_j - m)
  // (attention_kernels.cuh:560-664) -- pieces dealt round-robin to the warps, which meet in
  // shared memory (the ring is free by now) in warp order -- and leaves the counter at zero.
  __syncthreads();
  TRACE(25);
  if (threadIdx.x < 2 && pend_s[threadIdx.x][0]) {
    int* pd = pend_s[threadIdx.x];
    const int ticket = atomicAdd(p.counters + pd[1], 1);
    pd[5] = ticket == pd[2] - 1;
    if (pd[5]) atomicExch(p.counters + pd[1], 0);
  }
  __syncthreads();
  TRACE(26);
#pragma unroll 1
  for (int k = 0; k < 2; ++k) {
    if (!pend_s[k][0] || !pend_s[k][5]) continue;     // CTA-uniform
    __threadfence();                                   // the other CTAs' slots, fenced before their tickets
    constexpr int SF = slot_floats(D);
    const int c_first = pend_s[k][1], n = pend_s[k][2], b = pend_s[k][3], vh = pend_s[k][4];
    const float* const cta_slots = p.ws + (size_t)2 * p.num_workers * SF;
    // piece j: slot 2 c_first + 1 for j = 0, else 2 (c_first + j)
    auto slot_of = [&](int j) {
      return cta_slots + (size_t)(2 * (c_first + j) + (j == 0 ? 1 : 0)) * SF;
    };
    float4* acc_s = reinterpret_cast<float4*>(smem);          // [NW][MT][32]
    float* lsum_s = reinterpret_cast<float*>(smem + (size_t)NW * MT * 32 * sizeof(float4));  // [NW][8]
    // pass 1: per-head maximum of the pieces' references (lanes stride over pieces)
    float mx8[8];
#pragma unroll
    for (int h = 0; h < 8; ++h) mx8[h] = -INFINITY;
    for (int j = lane; j < n; j += 32) {
      const float4* m = reinterpret_cast<const float4*>(slot_of(j) + 8 * D);
      const float4 a = __ldcg(m), bb = __ldcg(m + 1);
      mx8[0] = fmaxf(mx8[0], a.x); mx8[1] = fmaxf(mx8[1], a.y);
      mx8[2] = fmaxf(mx8[2], a.z); mx8[3] = fmaxf(mx8[3], a.w);
      mx8[4] = fmaxf(mx8[4], bb.x); mx8[5] = fmaxf(mx8[5], bb.y);
      mx8[6] = fmaxf(mx8[6], bb.z); mx8[7] = fmaxf(mx8[7], bb.w);
    }
#pragma unroll
    for (int h = 0; h < 8; ++h) mx8[h] = warp_max(mx8[h]);
    float mx[2];
#pragma unroll
    for (int h = 0; h < 8; ++h) {   // select heads 2q4, 2q4+1 without dynamic register indexing
      if (h == 2 * q4) mx[0] = mx8[h];
      if (h == 2 * q4 + 1) mx[1] = mx8[h];
    }
    // pass 2: warp `warp` sums pieces warp, warp + NW, ... (whole slots)
    float4 acc[MT];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) acc[mt] = make_float4(0.f, 0.f, 0.f, 0.f);
    float ls[2] = {0.f, 0.f};
#pragma unroll 1
    for (int j = warp; j < n; j += NW) {
      const float* slot = slot_of(j);
      const float2 m = __ldcg(reinterpret_cast<const float2*>(slot + 8 * D + 2 * q4));
      const float2 l = __ldcg(reinterpret_cast<const float2*>(slot + 8 * D + 8 + 2 * q4));
      float4 v[MT];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) v[mt] = __ldcg(reinterpret_cast<const float4*>(slot) + mt * 32 + lane);
      const float f0 = piece_weight(m.x, mx[0]), f1 = piece_weight(m.y, mx[1]);
      ls[0] += l.x * f0;
      ls[1] += l.y * f1;
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        acc[mt].x += v[mt].x * f0; acc[mt].y += v[mt].y * f1;
        acc[mt].z += v[mt].z * f0; acc[mt].w += v[mt].w * f1;
      }
    }
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) acc_s[(warp * MT + mt) * 32 + lane] = acc[mt];
    if (r == 0) {
      lsum_s[warp * 8 + 2 * q4] = ls[0];
      lsum_s[warp * 8 + 2 * q4 + 1] = ls[1];
    }
    __syncthreads();
    // final: warp `warp` adds the warps' sums of m-tiles warp, warp + NW, ... in warp order
    float lt[2] = {0.f, 0.f};
#pragma unroll
    for (int g = 0; g < NW; ++g) {
      lt[0] += lsum_s[g * 8 + 2 * q4];
      lt[1] += lsum_s[g * 8 + 2 * q4 + 1];
    }
    const int kvh = vh / p.h_tiles;
    const int tile = vh - kvh * p.h_tiles;
    const int hd0 = kvh * p.group + tile * 8;
    const int hv = min(8, p.group - tile * 8);
    const float inv0 = __fdividef(1.f, lt[0] + 1e-6f);
    const float inv1 = __fdividef(1.f, lt[1] + 1e-6f);
    for (int mt = warp; mt < MT; mt += NW) {
      float o4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int g = 0; g < NW; ++g) {
        const float4 v = acc_s[(g * MT + mt) * 32 + lane];
        o4[0] += v.x; o4[1] += v.y; o4[2] += v.z; o4[3] += v.w;
      }
      store_out_tile<ST, D, DPERM>(reinterpret_cast<T*>(p.out) + ((int64_t)b * p.num_heads + hd0) * D,
                                   mt, o4, inv0, inv1, hv, r, q4);
    }
    if (p.lse != nullptr && warp == 0 && r == 0) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
        if (2 * q4 + h < hv)
          p.lse[b * p.lse_seq_stride + (hd0 + 2 * q4 + h) * p.lse_head_stride] =
              (mx[h] + __log2f(lt[h])) * kLn2;
    }
    __syncthreads();   // acc_s is reused by the CTA's other pending item
  }
  TRACE(27);
#undef TRACE
#undef TRACE_W
}

// ---------------------------------------------------------------------------------------
#if STREAM_PART == 6 || STREAM_PART == 7
#include "paged_attention_nhd.cuh"
#endif

constexpr int kMaxWarpsPerSm = 16;

template <int ST, bool FP8, int D, int NW, int STAGES, int PPS, int MINB, bool FLASH = false,
          int BSZ = 16, bool E5 = false>
int launch_stream_t(StreamParams& p, int num_sms, int64_t workspace_bytes, cudaStream_t stream) {
  constexpr int CES = FP8 ? 1 : 2;
  constexpr int SMEM = NW * STAGES * PPS * 2 * kBS * (FLASH ? flash_row_bytes(D * CES) : D * CES);
  auto kern = paged_attention_stream_kernel<ST, FP8, D, NW, STAGES, PPS, MINB, FLASH, BSZ, E5>;
  // per instantiation AND per device: the shared-memory opt-in is a per-device function
  // attribute (a process that drives several GPUs must set it on each)
  static int ctas_per_sm_of[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
    return set_error(B200PA_ERR_CUDA, "cudaGetDevice failed or device ordinal out of range");
  int& ctas_per_sm = ctas_per_sm_of[dev];
  if (ctas_per_sm == 0) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess)
      return set_error(B200PA_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    int n = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, NW * 32, SMEM);
    if (e != cudaSuccess || n <= 0)
      return set_error(B200PA_ERR_CUDA, "occupancy query failed: %s", cudaGetErrorString(e));
    ctas_per_sm = n * NW <= kMaxWarpsPerSm ? n : kMaxWarpsPerSm / NW;
  }
  const int ctas = num_sms * ctas_per_sm;
  p.num_workers = ctas * NW;
  p.nw = NW;
  if (ctas * (int)sizeof(int) > kCounterBytes)
    return set_error(B200PA_ERR_UNSUPPORTED, "stream kernel: %d CTAs exceed the workspace header", ctas);
  // header (ticket counters) + 2 slots per warp (pieces) + 2 per CTA (what the in-CTA merge
  // leaves for the cross-CTA merge)
  const int64_t need = kCounterBytes +
      (int64_t)2 * (p.num_workers + ctas) * slot_floats(D) * (int64_t)sizeof(float);
  if (workspace_bytes < need)
    return set_error(B200PA_ERR_INVALID, "workspace too small: %lld bytes, need %lld "
                     "(b200pa_paged_attention_workspace_bytes)", (long long)workspace_bytes,
                     (long long)need);
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ctas);
    cfg.blockDim = dim3(NW * 32);
    cfg.dynamicSmemBytes = SMEM;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
    if (e != cudaSuccess)
      return set_error(B200PA_ERR_CUDA, "stream kernel launch: %s", cudaGetErrorString(e));
  }
  count_launch();
  return check_launch("paged_attention (stream)");
}

}  // namespace

// The instantiations are spread over four translation units (this file compiled with
// -DSTREAM_PART=0..4, see build.py) so that they build in parallel:
//   0: entry point, support queries, the bf16 head-size-128 shapes (with the tuning variants)
//   1: token-major ("flash") layout     2: 32-token cache blocks     3: the other head sizes
//   4: fp8_e5m2 caches
#ifndef STREAM_PART
#define STREAM_PART 0
#endif
int launch_stream_flash(StreamParams& p, bool fp8, int scalar_type, int head_size, int block_size,
    int num_sms, int64_t workspace_bytes, cudaStream_t stream);
int launch_stream_bs32(StreamParams& p, bool fp8, int scalar_type, int head_size, int num_sms,
    int64_t workspace_bytes, cudaStream_t stream);
int launch_stream_other(StreamParams& p, bool fp8, int scalar_type, int head_size, int tune,
    int num_sms, int64_t workspace_bytes, cudaStream_t stream);
int launch_stream_e5m2(StreamParams& p, int scalar_type, int head_size, bool flash, int block_size,
    int num_sms, int64_t workspace_bytes, cudaStream_t stream);
int launch_stream_e5m2_wide(StreamParams& p, int scalar_type, int head_size, bool flash,
    int block_size, int num_sms, int64_t workspace_bytes, cudaStream_t stream);
// NHD caches with 2 / 4 / 8 kv heads per GPU (paged_attention_nhd.cuh); -1: not taken (the caller
// goes on with the stream kernel's FLASH path)
int launch_stream_block_xp(const StreamParams& p, bool fp8, int scalar_type, int head_size,
    int tune, void* workspace, int64_t workspace_bytes, int num_sms, cudaStream_t stream);
int launch_stream_block_nhd(const StreamParams& p, bool fp8, int scalar_type, int head_size,
    int tune, void* workspace, int64_t workspace_bytes, int num_sms, cudaStream_t stream);

#define STREAM_CALL(ST, FP8V, DV, NWV, SG, PP, MB, ...) \
  return launch_stream_t<ST, FP8V, DV, NWV, SG, PP, MB, ##__VA_ARGS__>(p, num_sms, workspace_bytes, stream)

#if STREAM_PART == 1
int launch_stream_flash(StreamParams& p, bool fp8, int scalar_type, int head_size, int block_size,
    int num_sms, int64_t workspace_bytes, cudaStream_t stream) {
    // one instantiation per (scalar_t, cache type, head size, block size): the shapes of the
    // x-packed kernel where the padded rows fit (fp8 D=128: 8 warps x 3 stages = 216 KiB)
#define STREAM_FLASH_BY_HEAD_SIZE(ST, BSV, NW128, S128, P128, S192, P192, NW256, S256, P256)             \
  switch (head_size) {                                                                       \
    case 32:  if (fp8) { STREAM_CALL(ST, true, 32, 4, 3, 2, 2, true, BSV); }  else { STREAM_CALL(ST, false, 32, 4, 3, 2, 2, true, BSV); }  \
    case 64:  if (fp8) { STREAM_CALL(ST, true, 64, 4, 3, 2, 2, true, BSV); }  else { STREAM_CALL(ST, false, 64, 4, 3, 2, 2, true, BSV); }  \
    case 80:  if (fp8) { STREAM_CALL(ST, true, 80, 4, 3, 2, 2, true, BSV); }  else { STREAM_CALL(ST, false, 80, (BSV == 16 ? 8 : 4), (BSV == 16 ? 2 : 3), 2, (BSV == 16 ? 1 : 2), true, BSV); }  \
    case 96:  if (fp8) { STREAM_CALL(ST, true, 96, 4, 3, 2, 2, true, BSV); }  else { STREAM_CALL(ST, false, 96, (BSV == 16 ? 8 : 4), (BSV == 16 ? 2 : 3), 2, 1, true, BSV); }  \
    case 112: if (fp8) { STREAM_CALL(ST, true, 112, 4, 3, 2, 2, true, BSV); } else { STREAM_CALL(ST, false, 112, (BSV == 16 ? 6 : 4), (BSV == 16 ? 2 : 3), 2, 1, true, BSV); } \
    case 128: if (fp8) { STREAM_CALL(ST, true, 128, 8, 3, 2, 1, true, BSV); } else { STREAM_CALL(ST, false, 128, NW128, S128, P128, 1, true, BSV); } \
    case 192: if (fp8) { STREAM_CALL(ST, true, 192, 4, 4, 2, 1, true, BSV); } else { STREAM_CALL(ST, false, 192, 4, S192, P192, 1, true, BSV); } \
    case 256: if (fp8) { STREAM_CALL(ST, true, 256, 4, 3, 2, 1, true, BSV); } else { STREAM_CALL(ST, false, 256, NW256, S256, P256, 1, true, BSV); } \
    default: return set_error(B200PA_ERR_UNSUPPORTED, "stream kernel: head size %d", head_size);  \
  }
    // 16-bit caches at head sizes >= 128: the launch shapes of the x-packed kernel for the same
    // block size (same spans, same merge order: bitwise the same output); a stage holds whole
    // cache blocks, so two tiles per stage with 32-token blocks
    if (block_size == 32) {
      if (scalar_type == B200PA_BF16) { STREAM_FLASH_BY_HEAD_SIZE(B200PA_BF16, 32, 4, 3, 2, 2, 2, 3, 2, 2) }
      else { STREAM_FLASH_BY_HEAD_SIZE(B200PA_F16, 32, 4, 3, 2, 2, 2, 3, 2, 2) }
    }
    if (scalar_type == B200PA_BF16) { STREAM_FLASH_BY_HEAD_SIZE(B200PA_BF16, 16, 8, 2, 1, 2, 1, 4, 2, 1) }
    else { STREAM_FLASH_BY_HEAD_SIZE(B200PA_F16, 16, 8, 2, 1, 2, 1, 4, 2, 1) }
#undef STREAM_FLASH_BY_HEAD_SIZE
}
#elif STREAM_PART == 2
int launch_stream_bs32(StreamParams& p, bool fp8, int scalar_type, int head_size, int num_sms,
    int64_t workspace_bytes, cudaStream_t stream) {
    // 32-token cache blocks: stages of two 16-token tiles = one block (PPS = 2 everywhere)
#define STREAM_BY_HEAD_SIZE_32(ST)                                                           \
  switch (head_size) {                                                                       \
    case 32:  if (fp8) { STREAM_CALL(ST, true, 32, 4, 3, 2, 2, false, 32); }  else { STREAM_CALL(ST, false, 32, 4, 3, 2, 2, false, 32); }  \
    case 64:  if (fp8) { STREAM_CALL(ST, true, 64, 4, 3, 2, 2, false, 32); }  else { STREAM_CALL(ST, false, 64, 4, 3, 2, 2, false, 32); }  \
    case 80:  if (fp8) { STREAM_CALL(ST, true, 80, 4, 3, 2, 2, false, 32); } else { STREAM_CALL(ST, false, 80, 4, 3, 2, 2, false, 32); }  \
    case 96:  if (fp8) { STREAM_CALL(ST, true, 96, 4, 3, 2, 2, false, 32); }  else { STREAM_CALL(ST, false, 96, 4, 3, 2, 1, false, 32); }  \
    case 112: if (fp8) { STREAM_CALL(ST, true, 112, 4, 3, 2, 2, false, 32); } else { STREAM_CALL(ST, false, 112, 4, 3, 2, 1, false, 32); } \
    case 128: if (fp8) { STREAM_CALL(ST, true, 128, 6, 2, 2, 2, false, 32); } else { STREAM_CALL(ST, false, 128, 4, 3, 2, 1, false, 32); } \
    case 192: if (fp8) { STREAM_CALL(ST, true, 192, 4, 4, 2, 1, false, 32); } else { STREAM_CALL(ST, false, 192, 4, 2, 2, 1, false, 32); } \
    case 256: if (fp8) { STREAM_CALL(ST, true, 256, 4, 3, 2, 1, false, 32); } else { STREAM_CALL(ST, false, 256, 3, 2, 2, 1, false, 32); } \
    default: return set_error(B200PA_ERR_UNSUPPORTED, "stream kernel: head size %d", head_size);  \
  }
    if (scalar_type == B200PA_BF16) { STREAM_BY_HEAD_SIZE_32(B200PA_BF16) }
    else { STREAM_BY_HEAD_SIZE_32(B200PA_F16) }
#undef STREAM_BY_HEAD_SIZE_32
}
#elif STREAM_PART == 3
int launch_stream_other(StreamParams& p, bool fp8, int scalar_type, int head_size, int tune,
    int num_sms, int64_t workspace_bytes, cudaStream_t stream) {
  // the other head sizes (and f16 at 128): one shape each.  192 / 256: 12-16 KiB stages, 4 warps.
#define STREAM_BY_HEAD_SIZE(ST)                                                              \
  switch (head_size) {                                                                       \
    case 32:  if (fp8) { STREAM_CALL(ST, true, 32, 4, 3, 2, 2); }  else { STREAM_CALL(ST, false, 32, 4, 3, 2, 2); }  \
    case 64:  if (fp8) { STREAM_CALL(ST, true, 64, 6, 2, 4, 2); }  else { STREAM_CALL(ST, false, 64, 4, 3, 2, 2); }  \
    case 80:  if (fp8) { STREAM_CALL(ST, true, 80, 4, 3, 2, 2); } else { STREAM_CALL(ST, false, 80, 8, 2, 2, 1); }  \
    case 96:  if (fp8) { STREAM_CALL(ST, true, 96, 4, 3, 2, 2); }  else { STREAM_CALL(ST, false, 96, 8, 2, 2, 1); }  \
    case 112: if (fp8) { STREAM_CALL(ST, true, 112, 6, 2, 2, 2); } else { STREAM_CALL(ST, false, 112, 6, 2, 2, 1); } \
    case 128: if (fp8) { STREAM_CALL(ST, true, 128, 6, 2, 2, 2); } else { STREAM_CALL(ST, false, 128, 8, 2, 1, 1); } \
    case 192: if (fp8) { STREAM_CALL(ST, true, 192, 8, 2, 2, 1); } else { STREAM_CALL(ST, false, 192, 4, 2, 1, 1); } \
    case 256: if (fp8) { STREAM_CALL(ST, true, 256, 6, 2, 2, 1); } else { STREAM_CALL(ST, false, 256, 4, 2, 1, 1); } \
    default: return set_error(B200PA_ERR_UNSUPPORTED, "stream kernel: head size %d", head_size);  \
  }
#if B200PA_TESTING
  // candidate shapes behind the tuning hook (bf16 queries): more warps, two stages -- what won at
  // head size 128 under the power cap (tools/shapebench.py --sustain)
  if (tune == 1 && scalar_type == B200PA_BF16) {
    switch (head_size) {
      case 64:  if (fp8) { STREAM_CALL(B200PA_BF16, true, 64, 4, 2, 4, 3); }  else { STREAM_CALL(B200PA_BF16, false, 64, 6, 2, 2, 2); }
      case 80:  if (fp8) { STREAM_CALL(B200PA_BF16, true, 80, 6, 2, 2, 2); }  else break;
      case 96:  if (fp8) { STREAM_CALL(B200PA_BF16, true, 96, 6, 2, 2, 2); }  else break;
      case 192: if (fp8) break; else { STREAM_CALL(B200PA_BF16, false, 192, 8, 2, 1, 1); }
      default: break;
    }
  }
  if (tune == 2 && scalar_type == B200PA_BF16) {
    switch (head_size) {
      case 64:  if (fp8) { STREAM_CALL(B200PA_BF16, true, 64, 8, 2, 4, 1); }  else { STREAM_CALL(B200PA_BF16, false, 64, 8, 3, 2, 1); }
      case 96:  if (fp8) { STREAM_CALL(B200PA_BF16, true, 96, 4, 2, 2, 3); }  else break;
      case 192: if (fp8) break; else { STREAM_CALL(B200PA_BF16, false, 192, 4, 2, 1, 1); }
      default: break;
    }
  }
#else
  (void)tune;
#endif
  if (scalar_type == B200PA_BF16) { STREAM_BY_HEAD_SIZE(B200PA_BF16) }
  else { STREAM_BY_HEAD_SIZE(B200PA_F16) }
#undef STREAM_BY_HEAD_SIZE
}
#elif STREAM_PART == 4 || STREAM_PART == 5
// fp8_e5m2 caches: the fp8 launch shapes of launch_stream_other / _bs32 / _flash.  Part 4: the
// x-packed layout, part 5: the token-major layout.
#define STREAM_E5M2_BY_HEAD_SIZE(ST, FL, BSV)                                     \
  switch (head_size) {                                                           \
    case 32:  STREAM_CALL(ST, true, 32, 4, 3, 2, 2, FL, BSV, true);              \
    case 64:  STREAM_CALL(ST, true, 64, ((FL || BSV == 32) ? 4 : 6), ((FL || BSV == 32) ? 3 : 2), ((FL || BSV == 32) ? 2 : 4), 2, FL, BSV, true); \
    case 80:  STREAM_CALL(ST, true, 80, 4, 3, 2, 2, FL, BSV, true);              \
    case 96:  STREAM_CALL(ST, true, 96, 4, 3, 2, 2, FL, BSV, true);              \
    case 112: STREAM_CALL(ST, true, 112, ((FL || BSV == 32) ? 4 : 6), ((FL || BSV == 32) ? 3 : 2), 2, 2, FL, BSV, true); \
    case 128: STREAM_CALL(ST, true, 128, (FL ? 8 : 6), (FL ? 3 : 2), 2, (FL ? 1 : 2), FL, BSV, true); \
    case 192: STREAM_CALL(ST, true, 192, ((FL || BSV == 32) ? 4 : 8), ((FL || BSV == 32) ? 4 : 2), 2, 1, FL, BSV, true); \
    case 256: STREAM_CALL(ST, true, 256, ((FL || BSV == 32) ? 4 : 6), ((FL || BSV == 32) ? 3 : 2), 2, 1, FL, BSV, true); \
    default: return set_error(B200PA_ERR_UNSUPPORTED,                            \
                              "stream kernel: head size %d with an fp8_e5m2 cache", head_size); \
  }
#if STREAM_PART == 4
int launch_stream_e5m2(StreamParams& p, int scalar_type, int head_size, bool flash, int block_size,
    int num_sms, int64_t workspace_bytes, cudaStream_t stream) {
  if (flash)
    return launch_stream_e5m2_wide(p, scalar_type, head_size, flash, block_size, num_sms,
                                   workspace_bytes, stream);
  if (block_size == 32) {
    if (scalar_type == B200PA_BF16) { STREAM_E5M2_BY_HEAD_SIZE(B200PA_BF16, false, 32) }
    else { STREAM_E5M2_BY_HEAD_SIZE(B200PA_F16, false, 32) }
  }
  if (scalar_type == B200PA_BF16) { STREAM_E5M2_BY_HEAD_SIZE(B200PA_BF16, false, 16) }
  else { STREAM_E5M2_BY_HEAD_SIZE(B200PA_F16, false, 16) }
}
#else
int launch_stream_e5m2_wide(StreamParams& p, int scalar_type, int head_size, bool flash,
    int block_size, int num_sms, int64_t workspace_bytes, cudaStream_t stream) {
  if (block_size == 32) {
    if (scalar_type == B200PA_BF16) { STREAM_E5M2_BY_HEAD_SIZE(B200PA_BF16, true, 32) }
    else { STREAM_E5M2_BY_HEAD_SIZE(B200PA_F16, true, 32) }
  }
  if (scalar_type == B200PA_BF16) { STREAM_E5M2_BY_HEAD_SIZE(B200PA_BF16, true, 16) }
  else { STREAM_E5M2_BY_HEAD_SIZE(B200PA_F16, true, 16) }
}
#endif
#undef STREAM_E5M2_BY_HEAD_SIZE
#elif STREAM_PART == 6 || STREAM_PART == 7
// part 6: the x-packed layout, part 7: the token-major NHD layout
static int launch_block_kernel(const StreamParams& sp, bool fp8, int scalar_type, int head_size,
    int tune, void* workspace, int64_t workspace_bytes, int num_sms, cudaStream_t stream) {
  constexpr bool XPV = STREAM_PART == 6;
  if (head_size != 128 || sp.group > 8 || sp.h_tiles != 1) return -1;
  if (sp.num_kv_heads % 2 != 0) return -1;     // head groups of 2 / 4 / 8
  if (sp.bs_vert_stride > 1) return -1;
  NhdParams p;
  p.out = sp.out;
  p.lse = sp.lse;
  p.lse_seq_stride = sp.lse_seq_stride;
  p.lse_head_stride = sp.lse_head_stride;
  p.ws = nullptr;
  p.counters = nullptr;
  p.q = sp.q;
  p.k_cache = sp.k_cache;
  p.v_cache = sp.v_cache;
  p.block_tables = sp.block_tables;
  p.seq_lens = sp.seq_lens;
  p.alibi_slopes = sp.alibi_slopes;
  p.k_scale = sp.k_scale;
  p.v_scale = sp.v_scale;
  p.new_key = sp.new_key;
  p.new_value = sp.new_value;
  p.slot_mapping = sp.slot_mapping;
  p.new_key_stride = sp.new_key_stride;
  p.new_value_stride = sp.new_value_stride;
  p.q_stride = sp.q_stride;
  p.kv_block_stride = sp.kv_block_stride;
  p.kv_token_stride = sp.kv_token_stride;
  p.kv_head_stride = sp.kv_head_stride;
  p.num_seqs = sp.num_seqs;
  p.num_heads = sp.num_heads;
  p.num_kv_heads = sp.num_kv_heads;
  p.group = sp.group;
  p.max_num_blocks_per_seq = sp.max_num_blocks_per_seq;
  p.scale = sp.scale;
  // launch shapes (kv heads per CTA x stages x blocks per stage x CTAs per SM), tune 20 + i:
  //   0: 8 x 3 x (2 | 1) x 1    1: 4 x 2 x (2 | 1) x 3    2: 4 x 3 x (2 | 1) x 2
  //   3: 2 x 2 x (2 | 1) x 6    4: 8 x 2 x (2 | 1) x 1        (fp8 | 16-bit blocks per stage)
  // Defaults by the sustained rate (profiles/r02m_kbench_block.jsonl, burst / last third of 3 s):
  // fp8 shape 1 -- 12 warps -- 6666 / 6242 GB/s on cfg4 (shape 0: 6884 / 5767; the per-warp
  // kernel 6387 / 6175); 16-bit shape 0, 6228 / 6460 on cfg2 (per-warp 6083 / 6292).
  int v = (tune >= 20 && tune <= 24) ? tune - 20 : (fp8 ? 1 : 0);
  const int nw_of[5] = {8, 4, 4, 2, 8};
  while (sp.num_kv_heads % nw_of[v] != 0) v = (v == 0 || v == 4) ? 1 : 3;
#define BLK_CALL(ST, FP8V, NWV, SG, BP, MB)                                                     \
  return launch_nhd_t<ST, FP8V, false, 128, NWV, SG, BP, XPV, MB>(p, workspace, workspace_bytes, \
                                                                 num_sms, stream)
#define BLK_BY_SHAPE(ST)                                                                        \
  switch (v) {                                                                                  \
    case 0: if (fp8) { BLK_CALL(ST, true, 8, 3, 2, 1); } else { BLK_CALL(ST, false, 8, 3, 1, 1); } \
    case 2: if (fp8) { BLK_CALL(ST, true, 4, 3, 2, 2); } else { BLK_CALL(ST, false, 4, 3, 1, 2); } \
    case 3: if (fp8) { BLK_CALL(ST, true, 2, 2, 2, 6); } else { BLK_CALL(ST, false, 2, 2, 1, 6); } \
    case 4: if (fp8) { BLK_CALL(ST, true, 8, 2, 2, 1); } else { BLK_CALL(ST, false, 8, 2, 1, 1); } \
    default: if (fp8) { BLK_CALL(ST, true, 4, 2, 2, 3); } else { BLK_CALL(ST, false, 4, 2, 1, 3); } \
  }
  if (scalar_type == B200PA_BF16) { BLK_BY_SHAPE(B200PA_BF16) }
  else { BLK_BY_SHAPE(B200PA_F16) }
#undef BLK_BY_SHAPE
#undef BLK_CALL
}
#if STREAM_PART == 6
int launch_stream_block_xp(const StreamParams& sp, bool fp8, int scalar_type, int head_size,
    int tune, void* workspace, int64_t workspace_bytes, int num_sms, cudaStream_t stream) {
  return launch_block_kernel(sp, fp8, scalar_type, head_size, tune, workspace, workspace_bytes,
                             num_sms, stream);
}
#else
int launch_stream_block_nhd(const StreamParams& sp, bool fp8, int scalar_type, int head_size,
    int tune, void* workspace, int64_t workspace_bytes, int num_sms, cudaStream_t stream) {
  return launch_block_kernel(sp, fp8, scalar_type, head_size, tune, workspace, workspace_bytes,
                             num_sms, stream);
}
#endif
#else
// Upper bound of the workspace any stream launch on a device with `num_sms` SMs needs.
int64_t stream_workspace_bytes(int num_sms, int head_size) {
  // warps per SM + (at most 4) CTAs per SM, two slots each
  return kCounterBytes + (int64_t)2 * num_sms * (kMaxWarpsPerSm + 4) * slot_floats(head_size) *
                             (int64_t)sizeof(float);
}

bool stream_supported(const void* query, const void* key_cache, const void* value_cache,
                      int num_heads, int num_kv_heads, int head_size, int block_size,
                      int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
                      int scalar_type, int kv_dtype, int blocksparse_vert_stride) {
  (void)blocksparse_vert_stride;   // block-sparse attention is served (masked, not skipped)
  if (block_size != 16 && block_size != 32) return false;
  // the reference's head sizes that are multiples of 16 (120 stays on the generic kernel)
  if (head_size != 32 && head_size != 64 && head_size != 80 && head_size != 96 &&
      head_size != 112 && head_size != 128 && head_size != 192 && head_size != 256)
    return false;
  if (scalar_type != B200PA_BF16 && scalar_type != B200PA_F16) return false;
  if (kv_dtype != B200PA_KV_AUTO && kv_dtype != B200PA_KV_FP8_E4M3 &&
      kv_dtype != B200PA_KV_FP8_E5M2)
    return false;
  if (num_kv_heads <= 0 || num_heads % num_kv_heads != 0) return false;
  const int ces = kv_dtype == B200PA_KV_AUTO ? 2 : 1;
  auto aligned = [](const void* p, uintptr_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; };
  // TMA bulk copies need 16-byte aligned sources; Q rows are read as 8-byte vectors
  if (!aligned(key_cache, 16) || !aligned(value_cache, 16) || !aligned(query, 8)) return false;
  if ((kv_block_stride * ces) % 16 != 0 || (kv_head_stride * ces) % 16 != 0) return false;
  if (kv_head_stride < (int64_t)head_size * block_size) return false;
  if (q_stride % 4 != 0) return false;
  return true;
}

// Token-major ("flash") caches: every token row of one head is one bulk copy.
bool stream_flash_supported(const void* query, const void* key_cache, const void* value_cache,
                            int num_heads, int num_kv_heads, int head_size, int block_size,
                            int64_t q_stride, int64_t block_stride, int64_t token_stride,
                            int64_t head_stride, int scalar_type, int kv_dtype) {
  if (block_size != 16 && block_size != 32) return false;
  if (head_size != 32 && head_size != 64 && head_size != 80 && head_size != 96 &&
      head_size != 112 && head_size != 128 && head_size != 192 && head_size != 256) return false;
  if (scalar_type != B200PA_BF16 && scalar_type != B200PA_F16) return false;
  if (kv_dtype != B200PA_KV_AUTO && kv_dtype != B200PA_KV_FP8_E4M3 &&
      kv_dtype != B200PA_KV_FP8_E5M2) return false;
  if (num_kv_heads <= 0 || num_heads % num_kv_heads != 0) return false;
  const int ces = kv_dtype == B200PA_KV_AUTO ? 2 : 1;
  auto aligned = [](const void* p, uintptr_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; };
  if (!aligned(key_cache, 16) || !aligned(value_cache, 16) || !aligned(query, 8)) return false;
  if ((block_stride * ces) % 16 != 0 || (token_stride * ces) % 16 != 0 ||
      (head_stride * ces) % 16 != 0) return false;
  if (token_stride <= 0 || head_stride < head_size) return false;
  if (q_stride % 4 != 0) return false;
  return true;
}

// The fused write reads the new key/value rows with 4-byte loads (pairs of 16-bit elements).
bool stream_fused_supported(const FusedWrite& f, int scalar_type) {
  if (scalar_type == B200PA_F32) return false;
  auto a4 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 3u) == 0; };
  return f.key && f.value && f.slot_mapping && a4(f.key) && a4(f.value) &&
         f.key_stride % 2 == 0 && f.value_stride % 2 == 0;
}

int launch_stream(void* out, float* lse, int64_t lse_seq_stride, int64_t lse_head_stride,
                  void* workspace, int64_t workspace_bytes, const void* query,
                  const void* key_cache, const void* value_cache, int num_seqs, int num_heads,
                  int num_kv_heads, int head_size, float scale, const int32_t* block_tables,
                  int max_num_blocks_per_seq, const int32_t* seq_lens, const float* alibi_slopes,
                  int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
                  int scalar_type, int kv_dtype, const float* k_scale, const float* v_scale,
                  const FusedWrite* fused, const BlockSparse* bsp, int64_t flash_token_stride,
                  int block_size, int num_sms, int tune, cudaStream_t stream) {
  // flash_token_stride > 0: token-major cache (kv_block_stride / flash_token_stride /
  // kv_head_stride are the strides of [block][token][head][D]); 0: the x-packed layout
  if (!workspace || (reinterpret_cast<uintptr_t>(workspace) & 15u))
    return set_error(B200PA_ERR_INVALID, "the stream kernel needs a 16-byte aligned workspace "
                     "(b200pa_paged_attention_workspace_bytes)");
  StreamParams p;
  p.out = out;
  p.lse = lse;
  p.lse_seq_stride = lse_seq_stride;
  p.lse_head_stride = lse_head_stride;
  p.counters = reinterpret_cast<int*>(workspace);
  p.trace = reinterpret_cast<unsigned long long*>(testing_trace_buffer());
  p.ws = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kCounterBytes);
  p.q = query;
  p.k_cache = key_cache;
  p.v_cache = value_cache;
  p.block_tables = block_tables;
  p.seq_lens = seq_lens;
  p.alibi_slopes = alibi_slopes;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.new_key = fused ? fused->key : nullptr;
  p.new_value = fused ? fused->value : nullptr;
  p.slot_mapping = fused ? fused->slot_mapping : nullptr;
  p.new_key_stride = fused ? fused->key_stride : 0;
  p.new_value_stride = fused ? fused->value_stride : 0;
  p.q_stride = q_stride;
  p.kv_block_stride = kv_block_stride;
  p.kv_head_stride = kv_head_stride;
  p.kv_token_stride = flash_token_stride;
  p.num_seqs = num_seqs;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.group = num_heads / num_kv_heads;
  p.h_tiles = (p.group + 7) / 8;
  p.max_num_blocks_per_seq = max_num_blocks_per_seq;
  const bool sparse = bsp != nullptr && bsp->enabled;
  p.bs_vert_stride = sparse ? bsp->vert_stride : 0;
  p.bs_local_blocks = sparse ? bsp->local_blocks : 0;
  p.bs_block_size = sparse ? bsp->block_size : 1;
  p.bs_head_sliding_step = sparse ? bsp->head_sliding_step : 0;
  p.bs_tp_rank = sparse ? bsp->tp_rank : 0;
  p.num_workers = 0;
  p.nw = 0;
  p.scale = scale;
  const bool fp8 = kv_dtype == B200PA_KV_FP8_E4M3;
  if (kv_dtype == B200PA_KV_FP8_E5M2)
    return launch_stream_e5m2(p, scalar_type, head_size, flash_token_stride > 0, block_size, num_sms,
                              workspace_bytes, stream);
  if (flash_token_stride > 0) {
    // several kv heads per GPU in NHD order (a token's heads are adjacent, the tokens of one head
    // are not): whole-block stages through tensor-map TMA, one warp per head
    // (paged_attention_nhd.cuh).  tune >= 100 (test hook) keeps the row copies of the FLASH path.
    if (block_size == 16 && kv_head_stride < flash_token_stride && tune < 100 &&
        (kv_dtype == B200PA_KV_AUTO || kv_dtype == B200PA_KV_FP8_E4M3)) {
      const int rc = launch_stream_block_nhd(p, fp8, scalar_type, head_size, tune, workspace,
                                             workspace_bytes, num_sms, stream);
      if (rc >= 0) return rc;
    }
    return launch_stream_flash(p, fp8, scalar_type, head_size, block_size, num_sms, workspace_bytes,
                               stream);
  }
  if (block_size == 32)
    return launch_stream_bs32(p, fp8, scalar_type, head_size, num_sms, workspace_bytes, stream);
  // all the kv heads of a block on this GPU and dense ([block][head] contiguous): whole-block
  // stages, one warp per head (paged_attention_nhd.cuh with 1-D bulk copies of all heads' K / V).
  // tune >= 100 (test hook) keeps the per-warp streams, launch shape tune - 100.
  if (tune < 100 && kv_head_stride == (int64_t)head_size * block_size &&
      (kv_dtype == B200PA_KV_AUTO || kv_dtype == B200PA_KV_FP8_E4M3)) {
    const int rc = launch_stream_block_xp(p, fp8, scalar_type, head_size, tune, workspace,
                                          workspace_bytes, num_sms, stream);
    if (rc >= 0) return rc;
  }
  // <ST, FP8, D, NW, STAGES, PPS, MINB>: fp8 stage = PPS x 4 KiB, 16-bit stage = PPS x 8 KiB (D=128).
  // Defaults (tune 0) by the SUSTAINED rate: seconds of back-to-back steps run under the 1 kW
  // power cap at 1.4-1.7 GHz, where 12 warps x 2 stages (fp8) hold 6.4-6.6 TB/s and the former
  // default, 8 warps x 3 stages (tune 8), falls from 6.9 TB/s (first 50 ms) to 5.7-5.9; for 16-bit
  // caches 8 warps x 2 stages beat 4 x 4 (tune 10) in both regimes
  // (profiles/r02e_powerprobe_tunes.jsonl, r02f_powerprobe_cfg{2,3,4}.jsonl -- tune numbers there
  // are those of commit e162116).
#define STREAM_TUNE_FP8(ST, DV)                                        \
  switch (tune) {                                                      \
    case 1: STREAM_CALL(ST, true, DV, 4, 4, 2, 1);                     \
    case 2: STREAM_CALL(ST, true, DV, 4, 6, 1, 2);                     \
    case 3: STREAM_CALL(ST, true, DV, 6, 4, 2, 1);                     \
    case 4: STREAM_CALL(ST, true, DV, 4, 3, 2, 2);                     \
    case 5: STREAM_CALL(ST, true, DV, 4, 2, 2, 3);                     \
    case 6: STREAM_CALL(ST, true, DV, 5, 4, 2, 1);                     \
    case 7: STREAM_CALL(ST, true, DV, 6, 3, 2, 1);                     \
    case 8: STREAM_CALL(ST, true, DV, 8, 3, 2, 1);                     \
    case 9: STREAM_CALL(ST, true, DV, 12, 2, 2, 1);                    \
    case 10: STREAM_CALL(ST, true, DV, 5, 2, 2, 2);                    \
    case 11: STREAM_CALL(ST, true, DV, 3, 2, 2, 4);                    \
    case 12: STREAM_CALL(ST, true, DV, 10, 2, 2, 1);                   \
    case 13: STREAM_CALL(ST, true, DV, 8, 2, 2, 1);                    \
    default: STREAM_CALL(ST, true, DV, 6, 2, 2, 2);                    \
  }
#define STREAM_TUNE_16(ST, DV)                                         \
  switch (tune) {                                                      \
    case 1: STREAM_CALL(ST, false, DV, 4, 2, 1, 3);                    \
    case 2: STREAM_CALL(ST, false, DV, 6, 2, 2, 1);                    \
    case 3: STREAM_CALL(ST, false, DV, 6, 4, 1, 1);                    \
    case 4: STREAM_CALL(ST, false, DV, 8, 3, 1, 1);                    \
    case 5: STREAM_CALL(ST, false, DV, 4, 3, 1, 2);                    \
    case 6: STREAM_CALL(ST, false, DV, 4, 6, 1, 1);                    \
    case 7: STREAM_CALL(ST, false, DV, 5, 4, 1, 1);                    \
    case 8: STREAM_CALL(ST, false, DV, 12, 2, 1, 1);                   \
    case 9: STREAM_CALL(ST, false, DV, 6, 2, 1, 2);                    \
    case 10: STREAM_CALL(ST, false, DV, 4, 4, 1, 1);                   \
    case 11: STREAM_CALL(ST, false, DV, 10, 2, 1, 1);                  \
    case 12: STREAM_CALL(ST, false, DV, 4, 2, 1, 2);                   \
    case 13: STREAM_CALL(ST, false, DV, 9, 2, 1, 1);                   \
    default: STREAM_CALL(ST, false, DV, 8, 2, 1, 1);                   \
  }
  tune = tune >= 100 ? tune - 100 : 0;     // launch shapes of the per-warp kernel: 100 + i
  if (scalar_type == B200PA_BF16 && head_size == 128) {
    if (fp8) { STREAM_TUNE_FP8(B200PA_BF16, 128) } else { STREAM_TUNE_16(B200PA_BF16, 128) }
  }
#undef STREAM_TUNE_FP8
#undef STREAM_TUNE_16
  return launch_stream_other(p, fp8, scalar_type, head_size, tune, num_sms, workspace_bytes, stream);
}
#endif
#undef STREAM_CALL

}  // namespace b200pa
