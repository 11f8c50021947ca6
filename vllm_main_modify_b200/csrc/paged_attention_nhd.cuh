// The whole-block decode kernel: paged decode attention when several kv heads live on the GPU,
// for the x-packed cache and for the token-major NHD cache -- included by
// paged_attention_stream.cu (inside its anonymous namespace: it shares the span arithmetic, the
// PTX wrappers, the fragment conventions and the output helpers of the per-warp stream kernel).
//
// Why a second work decomposition.  The stream kernel gives every WARP its own
// (sequence, head) stream: 2-4 KiB bulk copies, and in the NHD layout [block][token][head][D] --
// where a (block, head) is 16 rows of 128-256 bytes one `Hkv * D` apart -- 16-byte cp.async
// granules, the rows of the other heads of a block being fetched at some other time by some other
// warp (5.1 TB/s on cfg3 against 6.4 x-packed).  Here the unit is the cache block with ALL the
// heads of a head group:
//  * a CTA has one warp per kv head of a group of NW heads; an item is one (sequence, head group);
//    all stages (BPS blocks each) are flattened and cut into one span per CTA, and the CTA's
//    warps walk the span together, each computing its own head;
//  * x-packed: K (and V) of one block, NW adjacent heads, is ONE contiguous run -- one 1-D bulk
//    copy each (8-32 KiB), every warp finding its head's page at warp * PAGE_BYTES;
//  * NHD: a stage arrives through tensor-map TMA (cp.async.bulk.tensor.4d, SASS UTMALDG): the box
//    covers (128 bytes of D, 16 tokens, NW heads) of one block and the tensor map lists the
//    dimensions as (D, token, head, block), so the box lands TRANSPOSED to [head][token][128 B]:
//    every warp finds its head's 16 rows together, 128-byte swizzled (conflict-free ldmatrix, no
//    padding, no per-row copies);
//  * CTA-wide ring with full / empty mbarriers (warp 0 produces); a piece that does not cover
//    its whole sequence leaves one partial per head in the CTA's workspace slots and the CTAs of
//    an item elect the last one to arrive (ticket counter, as in the stream kernel) to merge them
//    -- every warp merges its own head, pieces in CTA order;
//  * the fused cache write of decode_step: the warp patches its head's row of the new token in
//    shared memory and stores it to the cache.
// Scope: block size 16, group size <= 8, head size 128 (NHD: head rows of 128 or 256 bytes), an
// even number of kv heads per GPU, auto / fp8_e4m3.  Everything else stays on the stream kernel.
#pragma once

struct NhdParams {
  void* out;
  float* lse;
  int64_t lse_seq_stride, lse_head_stride;
  float* ws;           // 2 * NW slots per CTA, slot_floats(D) floats each
  int* counters;       // workspace header: one ticket counter per CTA, zero between launches
  const void* q;
  const void* k_cache;        // x-packed layout: the bulk copies' source; both: the fused write's target
  const void* v_cache;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const float* alibi_slopes;
  const float* k_scale;
  const float* v_scale;
  // fused cache write (decode_step), new_key == nullptr: plain attention
  const void* new_key;
  const void* new_value;
  const int64_t* slot_mapping;
  int64_t new_key_stride, new_value_stride;           // elements per token
  int64_t q_stride;
  int64_t kv_block_stride, kv_token_stride, kv_head_stride;   // elements (token: NHD layout only)
  int num_seqs, num_heads, num_kv_heads, group;
  int max_num_blocks_per_seq;
  float scale;
};

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, int c0, int c1,
                                            int c2, int c3, uint32_t bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      ".L2::cache_hint [%0], [%1, {%2, %3, %4, %5}], [%6], %7;" ::"r"(dst),
      "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_plain(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// ST: scalar type of query / out.  FP8 / E5: cache format.  NW: kv heads per GPU = warps per CTA.
// BPS: cache blocks (16 tokens, all heads) per pipeline stage.
// XP: the cache is x-packed (key_cache [block][head][D/x][16][x], value_cache [block][head][D][16],
// heads dense): K and V of one block, ALL heads, are one contiguous run each -- one 1-D bulk copy
// each, and every warp finds its head's page at warp * PAGE_BYTES in the stream kernel's format.
// NW: kv heads per CTA (= warps); the kv heads of a GPU are cut into num_kv_heads / NW groups and
// an item is one (sequence, head group).  MINB: CTAs per SM the launch counts on.
template <int ST, bool FP8, bool E5, int D, int NW, int STAGES, int BPS, bool XP, int MINB>
__global__ void __launch_bounds__(NW * 32, MINB)
paged_attention_nhd_kernel(const __grid_constant__ CUtensorMap map_k,
                           const __grid_constant__ CUtensorMap map_v, const NhdParams p) {
  using T = typename Scalar<ST>::type;
  constexpr bool MMA_BF16 = (!FP8) && (ST == B200PA_BF16);
  constexpr int CES = FP8 ? 1 : 2;
  constexpr int ROW = D * CES;                   // bytes of one token row of one head
  constexpr int NCH = XP ? 1 : ROW / 128;        // NHD: 128-byte boxes per row
  static_assert(XP || (ROW % 128 == 0 && NCH >= 1 && NCH <= 2), "head rows of 128 or 256 bytes");
  constexpr int BOX_BYTES = 128 * kBS * NW;      // NHD: one box, [head][token][128 B]
  constexpr int HEAD_BYTES = 128 * kBS;          // NHD: one head's 16 rows inside a box
  constexpr int PAGE_BYTES = ROW * kBS;          // x-packed: one head's K (or V) page
  constexpr int KBLK = ROW * kBS * NW;           // K (or V) of one block, all heads
  constexpr int STAGE_BYTES = BPS * 2 * KBLK;
  constexpr int KCS = kBS * 16;                  // x-packed K: bytes between 16-byte chunks
  constexpr int VRS = kBS * CES;                 // x-packed V: bytes per head-element row
  constexpr int KSTEPS = D / 16;
  constexpr int MT = D / 16;
  constexpr bool DPERM = FP8 && !XP;             // see store_out_tile
  constexpr float OFF = MMA_BF16 ? 0.f : 6.f;
  constexpr float CAP = MMA_BF16 ? 8.f : 14.f;
  constexpr int SF = slot_floats(D);

  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[2 * STAGES];   // full[STAGES], empty[STAGES]
  __shared__ int pend_s[2][6];    // valid, first CTA of the item, CTAs of the item, b, last?, head group

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = lane >> 2, q4 = lane & 3;
  const int mat = lane >> 3, mi = lane & 7;
  asm volatile("griddepcontrol.wait;" ::: "memory");

  // ---------------------------------------------------------------- the CTA's span
  const int groups = p.num_kv_heads / NW;
  const long long Ttot = total_stages<BPS>(p.seq_lens, p.num_seqs, lane) * groups;
  const long long W = min((long long)gridDim.x, Ttot);
  const long long w = blockIdx.x;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (threadIdx.x < 2) pend_s[threadIdx.x][0] = 0;
  // 128-byte swizzle wants 1 KiB aligned boxes (the launch asks for 1 KiB of slack)
  const uint32_t ring_u32 = (smem_u32(smem) + 1023u) & ~1023u;
  const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[STAGES]);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, NW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  int mflags = 0;            // bit 0: head piece left in slot 0, bit 2: tail piece in slot 1
  int mh_b = 0, mt_b = 0, mh_ns = 0, mt_ns = 0, mh_vh = 0, mt_vh = 0;
  long long mh_ib = 0, mt_ib = 0;
  float* const my_slots = p.ws + (size_t)2 * blockIdx.x * NW * SF;   // [kind][warp]
  if (w < W) {
  const long long u0 = (w * Ttot) / W, u1 = ((w + 1) * Ttot) / W;
  Cursor pc;
  {
    const Loc loc = locate_stage<BPS>(p.seq_lens, p.num_seqs, groups, u0, lane);
    pc.b = loc.b; pc.vh = loc.vh; pc.len = loc.len; pc.ns = loc.ns;
    pc.left = (int)(u1 - u0);
    cursor_open(pc, loc.st);
  }
  Cursor cc = pc;
  int piece_begin = cc.st;
  int it = 0;
  const uint64_t policy = l2_evict_first_policy();

  // ---------------------------------------------------------------- producer (warp 0)
  int bt_reg = 0, bt_b = -1, bt_win = -1;
  bool producing = true;
  int ip = 0;
  auto produce = [&]() {   // every lane of warp 0
    if (pc.st == pc.st_end) {
      if (pc.left == 0) { producing = false; return; }
      cursor_next_piece<BPS>(pc, p.seq_lens, groups);
    }
    const int blk0 = pc.st * BPS;
    const int nblk = min(BPS, (pc.len + kBS - 1) / kBS - blk0);
    ++pc.st;
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
    if (lane == 0) {
      const int s = ip % STAGES;
      if (ip >= STAGES) mbar_wait(empty0 + 8 * s, ((ip / STAGES) - 1) & 1);
      const uint32_t bar = full0 + 8 * s;
      const uint32_t dst = ring_u32 + s * STAGE_BYTES;
      mbar_expect_tx(bar, nblk * 2 * KBLK);
#pragma unroll
      for (int g = 0; g < BPS; ++g) {
        if (g < nblk) {
          if constexpr (XP) {
            const int64_t boff = ((int64_t)phys[g] * p.kv_block_stride +
                                  (int64_t)pc.vh * NW * p.kv_head_stride) * CES;
            tma_load_1d(dst + g * 2 * KBLK, reinterpret_cast<const uint8_t*>(p.k_cache) + boff, KBLK,
                        bar, policy);
            tma_load_1d(dst + g * 2 * KBLK + KBLK,
                        reinterpret_cast<const uint8_t*>(p.v_cache) + boff, KBLK, bar, policy);
            continue;
          }
#pragma unroll
          for (int h = 0; h < NCH; ++h) {
            // box h of the block: elements [h * 128 / CES, + 128 / CES) of D, 16 tokens, NW heads
            tma_load_4d(dst + g * 2 * KBLK + h * BOX_BYTES, &map_k, h * (128 / CES), 0, pc.vh * NW,
                        phys[g], bar, policy);
            tma_load_4d(dst + g * 2 * KBLK + KBLK + h * BOX_BYTES, &map_v, h * (128 / CES), 0,
                        pc.vh * NW, phys[g], bar, policy);
          }
        }
      }
    }
    __syncwarp();
    ++ip;
  };
  if (warp == 0) {
#pragma unroll 1
    for (int i = 0; i < STAGES - 1 && producing; ++i) produce();
  }

  // ---------------------------------------------------------------- consumer state (head = warp)
  uint32_t qb[KSTEPS][2];
  float o[MT][4];
  float msub[2], lsum[2], slope_l2[2];
  const bool has_alibi = p.alibi_slopes != nullptr;
  int head0 = 0;                               // first query head of this warp's kv head
  const int heads_valid = p.group;            // <= 8
  float sl2 = p.scale * kLog2e, out_scale = 1.f;
  if constexpr (FP8) {
    sl2 *= *p.k_scale;
    out_scale = *p.v_scale;
  }
  // NHD: ldmatrix lane addressing inside a head's [16 tokens][128 B] tile, 128-byte swizzle:
  // chunk c of row t sits at t * 128 + ((c ^ (t & 7)) << 4); every lane's row has (t & 7) == mi.
  // K x4 = {tok 0-7, chunk c | tok 8-15, chunk c | tok 0-7, chunk c+1 | tok 8-15, chunk c+1}
  // V x4 (transposing): fp8 like K; 16-bit {tok 0-7, ch c | tok 0-7, ch c+1 | tok 8-15, ch c | ..}
  // x-packed: the stream kernel's addressing inside the head's page (see there).
  uint32_t k_row_off, v_row_off;
  int k_ch = 0, v_ch = 0, tokA, tokB;
  if constexpr (XP) {
    if constexpr (FP8) {
      const int sig0 = 4 * (mi >> 1) + (mi & 1) + ((mi >= 4) ? 2 : 0);
      const int tok = (mat & 1) ? (sig0 ^ 2) : sig0;
      k_row_off = warp * PAGE_BYTES + (mat >> 1) * KCS + tok * 16;
      v_row_off = warp * PAGE_BYTES + lane * VRS;
      tokA = 4 * (r >> 1) + (r & 1) + ((r >= 4) ? 2 : 0);
      tokB = tokA ^ 2;
    } else {
      k_row_off = warp * PAGE_BYTES + (mat >> 1) * KCS + ((mat & 1) * 8 + mi) * 16;
      v_row_off = warp * PAGE_BYTES + ((mat & 1) * 8 + mi) * VRS + (mat >> 1) * 16;
      tokA = r;
      tokB = r + 8;
    }
  } else {
    k_row_off = warp * HEAD_BYTES + ((mat & 1) * 8 + mi) * 128;
    k_ch = mat >> 1;
    v_row_off = FP8 ? k_row_off : warp * HEAD_BYTES + ((mat >> 1) * 8 + mi) * 128;
    v_ch = FP8 ? (mat >> 1) : (mat & 1);
    tokA = r;
    tokB = r + 8;
  }
  auto swz = [&](int cg) {   // NHD: chunk cg (0 .. ROW/16) of this lane's row -> byte offset in the block
    return (cg >> 3) * BOX_BYTES + (((cg & 7) ^ mi) << 4);
  };
  // fused cache write: this lane's pairs of head elements (2(lane + 32j), +1) of the new token
  constexpr int NP = (D / 2 + 31) / 32;
  uint32_t nt_k[NP], nt_v[NP];
  int64_t nt_slot = -1;
  float ks_raw = 1.f;
  if constexpr (FP8) ks_raw = *p.k_scale;

  auto begin_piece = [&]() {
    piece_begin = cc.st;
    const int kvh = cc.vh * NW + warp;
    head0 = kvh * p.group;
    if (p.new_key != nullptr) {
      nt_slot = -1;
      if (cc.st_end == cc.ns) {     // this piece reaches the end of the sequence
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
    const T* qrow = reinterpret_cast<const T*>(p.q) + (int64_t)cc.b * p.q_stride +
                    (int64_t)(head0 + r) * D;
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) {
      uint32_t v0 = 0, v1 = 0;
      if (r < heads_valid) {
        if constexpr (FP8) {
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
      const float inv0 = out_scale * __fdividef(1.f, lsum[0] + 1e-6f);
      const float inv1 = out_scale * __fdividef(1.f, lsum[1] + 1e-6f);
      store_out<ST, D, DPERM>(reinterpret_cast<T*>(p.out) + ((int64_t)cc.b * p.num_heads + head0) * D,
                              o, inv0, inv1, heads_valid, r, q4);
      if (p.lse != nullptr && r == 0) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
          if (2 * q4 + h < heads_valid)
            p.lse[cc.b * p.lse_seq_stride + (head0 + 2 * q4 + h) * p.lse_head_stride] =
                (msub[h] + __log2f(lsum[h])) * kLn2;
      }
    } else {
      const long long ib = u0 + it - cc.st_end;   // the sequence's first flattened stage
      if (from_start) { mflags |= 4; mt_b = cc.b; mt_vh = cc.vh; mt_ns = cc.ns; mt_ib = ib; }
      else { mflags |= 1; mh_b = cc.b; mh_vh = cc.vh; mh_ns = cc.ns; mh_ib = ib; }
      float* slot = my_slots + (size_t)((from_start ? NW : 0) + warp) * SF;
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
        reinterpret_cast<float4*>(slot)[mt * 32 + lane] = make_float4(
            o[mt][0] * out_scale, o[mt][1] * out_scale, o[mt][2] * out_scale, o[mt][3] * out_scale);
      if (r == 0) {
        *reinterpret_cast<float2*>(slot + 8 * D + 2 * q4) = make_float2(msub[0], msub[1]);
        *reinterpret_cast<float2*>(slot + 8 * D + 8 + 2 * q4) = make_float2(lsum[0], lsum[1]);
      }
    }
  };

  begin_piece();

  // ---------------------------------------------------------------- main loop
#pragma unroll 1
  for (;; ++it) {
    if (cc.st == cc.st_end) {
      end_piece();
      if (cc.left == 0) break;
      cursor_next_piece<BPS>(cc, p.seq_lens, groups);
      begin_piece();
    }
    if (warp == 0 && producing) produce();   // the stage the previous iteration released
    const int page0 = cc.st * BPS;
    const int npg = min(BPS, (cc.len + kBS - 1) / kBS - page0);
    ++cc.st;
    const int s = it % STAGES;
    const uint32_t sbase = ring_u32 + s * STAGE_BYTES;
    mbar_wait(full0 + 8 * s, (it / STAGES) & 1);

    // ---- fused cache write: the stage that ends the sequence holds position len - 1, whose K/V
    // come from this step's key / value rows (quantised exactly as reshape_and_cache does), not
    // from the bytes the copy found in the cache.  Every warp patches its own head only.
    if (p.new_key != nullptr && cc.st == cc.ns && nt_slot >= 0) {
      const int t = cc.len - 1 - page0 * kBS;
      const int g = t >> 4, tt = t & (kBS - 1);
      uint8_t* const stage = smem + (ring_u32 - smem_u32(smem)) + s * STAGE_BYTES + g * 2 * KBLK;
      const int64_t blk = nt_slot / kBS;
      const int off = (int)(nt_slot % kBS);
      uint8_t* const kg = reinterpret_cast<uint8_t*>(const_cast<void*>(p.k_cache)) +
                          (blk * p.kv_block_stride + (int64_t)(cc.vh * NW + warp) * p.kv_head_stride) * CES;
      uint8_t* const vg = reinterpret_cast<uint8_t*>(const_cast<void*>(p.v_cache)) +
                          (blk * p.kv_block_stride + (int64_t)(cc.vh * NW + warp) * p.kv_head_stride) * CES;
      auto put = [&](uint8_t* dst, uint32_t wv) {
        if constexpr (FP8) *reinterpret_cast<uint16_t*>(dst) = (uint16_t)wv;
        else *reinterpret_cast<uint32_t*>(dst) = wv;
      };
#pragma unroll
      for (int j = 0; j < NP; ++j) {
        const int e0 = 2 * (lane + 32 * j);
        if (e0 < D) {
          uint32_t kw = nt_k[j], vw = nt_v[j];
          if constexpr (FP8) {
            const T* ke = reinterpret_cast<const T*>(&nt_k[j]);
            const T* ve = reinterpret_cast<const T*>(&nt_v[j]);
            kw = fp8_quant2<(E5 ? B200PA_KV_FP8_E5M2 : B200PA_KV_FP8_E4M3)>(
                Scalar<ST>::to_float(ke[0]), Scalar<ST>::to_float(ke[1]), ks_raw);
            vw = fp8_quant2<(E5 ? B200PA_KV_FP8_E5M2 : B200PA_KV_FP8_E4M3)>(
                Scalar<ST>::to_float(ve[0]), Scalar<ST>::to_float(ve[1]), out_scale);
          }
          if constexpr (XP) {
            constexpr int X = 16 / CES;
            const int kchunk = (e0 / X) * KCS + (e0 % X) * CES;
            put(stage + warp * PAGE_BYTES + kchunk + tt * 16, kw);
            put(kg + kchunk + off * 16, kw);
#pragma unroll
            for (int i = 0; i < 2; ++i) {   // V [D][16]: one element per head-element row
              if constexpr (FP8) {
                const uint8_t b8 = (uint8_t)(vw >> (8 * i));
                stage[KBLK + warp * PAGE_BYTES + (e0 + i) * kBS + tt] = b8;
                vg[(e0 + i) * kBS + off] = b8;
              } else {
                const uint16_t v16 = (uint16_t)(vw >> (16 * i));
                reinterpret_cast<uint16_t*>(stage + KBLK + warp * PAGE_BYTES)[(e0 + i) * kBS + tt] = v16;
                reinterpret_cast<uint16_t*>(vg)[(e0 + i) * kBS + off] = v16;
              }
            }
          } else {
            // token-major rows: chunk (e0 * CES) / 16 of row tt, swizzled in shared memory
            const int byte = e0 * CES, cg = byte >> 4;
            const int soff = (cg >> 3) * BOX_BYTES + warp * HEAD_BYTES + tt * 128 +
                             (((cg & 7) ^ (tt & 7)) << 4) + (byte & 15);
            put(stage + soff, kw);
            put(stage + KBLK + soff, vw);
            put(kg + (int64_t)off * p.kv_token_stride * CES + byte, kw);
            put(vg + (int64_t)off * p.kv_token_stride * CES + byte, vw);
          }
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
    }

    // ---- S^T = K Q^T
    float sc[BPS][4];
#pragma unroll
    for (int g = 0; g < BPS; ++g) {
      sc[g][0] = sc[g][1] = sc[g][2] = sc[g][3] = 0.f;
      if (g < npg) {
        const uint32_t kbase = sbase + g * 2 * KBLK + k_row_off;
        float sb[4] = {0.f, 0.f, 0.f, 0.f};
        if constexpr (XP && FP8) {
#pragma unroll
          for (int c = 0; c < KSTEPS; c += 2) {
            uint32_t w4[4], a[4];
            ldmatrix_x4(w4, kbase + c * KCS);
            fp8x4_to_f16x2x2<E5>(w4[0], a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(w4[1], a[1], a[3]);
            mma16816<false>(sc[g], a, qb[c][0], qb[c][1]);
            fp8x4_to_f16x2x2<E5>(w4[2], a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(w4[3], a[1], a[3]);
            mma16816<false>(sb, a, qb[c + 1][0], qb[c + 1][1]);
          }
        } else if constexpr (XP) {
#pragma unroll
          for (int ks = 0; ks < KSTEPS; ++ks) {
            uint32_t a[4];
            ldmatrix_x4(a, kbase + ks * 2 * KCS);
            if (ks & 1) mma16816<MMA_BF16>(sb, a, qb[ks][0], qb[ks][1]);
            else mma16816<MMA_BF16>(sc[g], a, qb[ks][0], qb[ks][1]);
          }
        } else if constexpr (FP8) {
#pragma unroll
          for (int c = 0; c < KSTEPS; c += 2) {
            uint32_t w4[4], a[4];
            ldmatrix_x4(w4, kbase + swz(c + k_ch));
            fp8x4_to_f16x2x2<E5>(w4[0], a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(w4[1], a[1], a[3]);
            mma16816<false>(sc[g], a, qb[c][0], qb[c][1]);
            fp8x4_to_f16x2x2<E5>(w4[2], a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(w4[3], a[1], a[3]);
            mma16816<false>(sb, a, qb[c + 1][0], qb[c + 1][1]);
          }
        } else {
#pragma unroll
          for (int ks = 0; ks < KSTEPS; ++ks) {
            uint32_t a[4];
            ldmatrix_x4(a, kbase + swz(2 * ks + k_ch));
            if (ks & 1) mma16816<MMA_BF16>(sb, a, qb[ks][0], qb[ks][1]);
            else mma16816<MMA_BF16>(sc[g], a, qb[ks][0], qb[ks][1]);
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) sc[g][j] += sb[j];
      }
    }

    // ---- online softmax (the stream kernel's: log2 domain, lazily moved reference)
    const int tok0 = page0 * kBS;
    float d[BPS][4];
    auto logits = [&](float neg0, float neg1) {
#pragma unroll
      for (int g = 0; g < BPS; ++g) {
        const int nv = cc.len - (tok0 + g * kBS);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int tk = (j & 2) ? tokB : tokA;
          float t = fmaf(sc[g][j], sl2, (j & 1) ? neg1 : neg0);
          if (has_alibi) t += slope_l2[j & 1] * (float)(tok0 + g * kBS + tk - cc.len + 1);
          d[g][j] = (g < npg && tk < nv) ? t : -INFINITY;
        }
      }
    };
    logits(-msub[0], -msub[1]);
    bool over = false;
#pragma unroll
    for (int g = 0; g < BPS; ++g)
#pragma unroll
      for (int j = 0; j < 4; ++j) over |= d[g][j] > CAP;
    if (__any_sync(kFull, over)) {
      logits(0.f, 0.f);
      float alpha[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float mx = -INFINITY;
#pragma unroll
        for (int g = 0; g < BPS; ++g) mx = fmaxf(mx, fmaxf(d[g][h], d[g][2 + h]));
        mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, 4));
        mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, 8));
        mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, 16));
        alpha[h] = 1.f;
        if (mx > msub[h] + CAP) {
          const float nm = mx - OFF;
          alpha[h] = ex2(msub[h] - nm);
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
    uint32_t pb[BPS][2];
#pragma unroll
    for (int g = 0; g < BPS; ++g) {
      float pr[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) pr[j] = ex2(d[g][j]);
      lsum[0] += pr[0] + pr[2];
      lsum[1] += pr[1] + pr[3];
      const uint32_t ta = movmatrix_trans(pack2<MMA_BF16>(pr[0], pr[1]));
      const uint32_t tb = movmatrix_trans(pack2<MMA_BF16>(pr[2], pr[3]));
      if constexpr (XP && FP8) {     // see the stream kernel: b0 / b1 swapped for q >= 2
        pb[g][0] = (q4 >= 2) ? tb : ta;
        pb[g][1] = (q4 >= 2) ? ta : tb;
      } else {
        pb[g][0] = ta;
        pb[g][1] = tb;
      }
    }

    // ---- O^T += V^T P^T
#pragma unroll
    for (int g = 0; g < BPS; ++g) {
      if (g < npg) {
        const uint32_t vbase = sbase + g * 2 * KBLK + KBLK + v_row_off;
        // tail page: V rows past the context may hold NaN bit patterns (zeroed like the reference)
        const int nv = cc.len - (tok0 + g * kBS);
        uint32_t m01 = 0xffffffffu, m23 = 0xffffffffu;
        if (nv < kBS) {
          const int t01 = (XP && FP8) ? 4 * q4 : 2 * q4;
          const int t23 = (XP && FP8) ? 4 * q4 + 2 : 2 * q4 + 8;
          m01 = (t01 < nv ? 0x0000ffffu : 0u) | (t01 + 1 < nv ? 0xffff0000u : 0u);
          m23 = (t23 < nv ? 0x0000ffffu : 0u) | (t23 + 1 < nv ? 0xffff0000u : 0u);
        }
        if constexpr (XP && FP8) {
#pragma unroll
          for (int gg = 0; gg < D / 32; ++gg) {
            uint32_t w4[4], a[4];
            ldmatrix_x4(w4, vbase + gg * (32 * VRS));
            fp8x4_to_f16x2x2<E5>(w4[0], a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(w4[1], a[1], a[3]);
            a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23;
            mma16816<false>(o[2 * gg], a, pb[g][0], pb[g][1]);
            fp8x4_to_f16x2x2<E5>(w4[2], a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(w4[3], a[1], a[3]);
            a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23;
            mma16816<false>(o[2 * gg + 1], a, pb[g][0], pb[g][1]);
          }
        } else if constexpr (XP) {
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            uint32_t a[4];
            ldmatrix_x4(a, vbase + mt * (16 * VRS));
            a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23;
            mma16816<MMA_BF16>(o[mt], a, pb[g][0], pb[g][1]);
          }
        } else if constexpr (FP8) {
#pragma unroll
          for (int gg = 0; gg < D / 32; ++gg) {
            uint32_t w4[4], a[4];
            ldmatrix_x4_trans(w4, vbase + swz(2 * gg + v_ch));
            fp8x4_to_f16x2x2<E5>(prmt<0x6420>(w4[0], w4[1]), a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(prmt<0x7531>(w4[0], w4[1]), a[1], a[3]);
            a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23;
            mma16816<false>(o[2 * gg], a, pb[g][0], pb[g][1]);
            fp8x4_to_f16x2x2<E5>(prmt<0x6420>(w4[2], w4[3]), a[0], a[2]);
            fp8x4_to_f16x2x2<E5>(prmt<0x7531>(w4[2], w4[3]), a[1], a[3]);
            a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23;
            mma16816<false>(o[2 * gg + 1], a, pb[g][0], pb[g][1]);
          }
        } else {
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            uint32_t a[4];
            ldmatrix_x4_trans(a, vbase + swz(2 * mt + v_ch));
            a[0] &= m01; a[1] &= m01; a[2] &= m23; a[3] &= m23;
            mma16816<MMA_BF16>(o[mt], a, pb[g][0], pb[g][1]);
          }
        }
      }
    }

    // ---- this warp is done with the stage
    __syncwarp();
    if (lane == 0) mbar_arrive_plain(empty0 + 8 * s);
  }
  }  // w < W

  // empty sequences: zeros (the reference's v1 kernel: no keys, zero accumulators)
  for (long long b = (long long)blockIdx.x * NW + warp; b < p.num_seqs; b += (long long)gridDim.x * NW) {
    if (p.seq_lens[b] <= 0) {
      T* row = reinterpret_cast<T*>(p.out) + b * p.num_heads * D;
      for (int i = lane; i < p.num_heads * D; i += 32) row[i] = Scalar<ST>::from_float(0.f);
      if (p.lse != nullptr)
        for (int i = lane; i < p.num_heads; i += 32)
          p.lse[b * p.lse_seq_stride + i * p.lse_head_stride] = -INFINITY;
    }
  }

  // ---------------------------------------------------------------- cross-CTA merge
  // A sequence that CTA boundaries cut has one piece per CTA it touches and per head: slot
  // [kind 1][head] of its first CTA, [kind 0][head] of every later one.  Each of those CTAs takes
  // a ticket from the counter of the sequence's first CTA; the last to arrive sums the pieces
  // with the reference's weights l_j * 2^(m_j - m) (attention_kernels.cuh:560-664), every warp
  // its own head, pieces in CTA order (deterministic), and leaves the counter at zero.
  __threadfence();     // this warp's slots are visible device-wide before the CTA takes tickets
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      if (!(mflags & (k ? 4 : 1))) continue;
      const long long ib = k ? mt_ib : mh_ib, ie = ib + (k ? mt_ns : mh_ns);
      const int c_first = (int)(((ib + 1) * W - 1) / Ttot);
      const int c_last = (int)((ie * W - 1) / Ttot);
      const int n = c_last - c_first + 1;
      const int ticket = atomicAdd(p.counters + c_first, 1);
      const int last = ticket == n - 1;
      if (last) atomicExch(p.counters + c_first, 0);
      pend_s[k][1] = c_first;
      pend_s[k][2] = n;
      pend_s[k][3] = k ? mt_b : mh_b;
      pend_s[k][4] = last;
      pend_s[k][5] = k ? mt_vh : mh_vh;
      pend_s[k][0] = 1;
    }
  }
  __syncthreads();
#pragma unroll 1
  for (int k = 0; k < 2; ++k) {
    if (!pend_s[k][0] || !pend_s[k][4]) continue;      // CTA-uniform
    __threadfence();
    const int c_first = pend_s[k][1], n = pend_s[k][2], b = pend_s[k][3];
    auto slot_of = [&](int j) {
      return p.ws + ((size_t)2 * (c_first + j) * NW + (j == 0 ? NW : 0) + warp) * SF;
    };
    // pass 1: per-head maximum of the pieces' references
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
    for (int h = 0; h < 8; ++h) {
      if (h == 2 * q4) mx[0] = mx8[h];
      if (h == 2 * q4 + 1) mx[1] = mx8[h];
    }
    // pass 2: this warp's head, pieces in CTA order
    float4 acc[MT];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) acc[mt] = make_float4(0.f, 0.f, 0.f, 0.f);
    float ls[2] = {0.f, 0.f};
#pragma unroll 1
    for (int j = 0; j < n; ++j) {
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
    const float inv0 = __fdividef(1.f, ls[0] + 1e-6f), inv1 = __fdividef(1.f, ls[1] + 1e-6f);
    const int hd0 = (pend_s[k][5] * NW + warp) * p.group, hv = p.group;
    T* dst = reinterpret_cast<T*>(p.out) + ((int64_t)b * p.num_heads + hd0) * D;
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      const float o4[4] = {acc[mt].x, acc[mt].y, acc[mt].z, acc[mt].w};
      store_out_tile<ST, D, DPERM>(dst, mt, o4, inv0, inv1, hv, r, q4);
    }
    if (p.lse != nullptr && r == 0) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
        if (2 * q4 + h < hv)
          p.lse[b * p.lse_seq_stride + (hd0 + 2 * q4 + h) * p.lse_head_stride] =
              (mx[h] + __log2f(ls[h])) * kLn2;
    }
  }
}

// cuTensorMapEncodeTiled through the runtime (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) ==
            cudaSuccess && qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// Tensor map of one cache: dimensions listed as (D, token, head, block) -- whatever their order
// in memory (NHD or HND: the strides say) -- and a box of (128 bytes of D, 16 tokens, NW heads),
// which therefore lands in shared memory as [head][token][128 B], 128-byte swizzled.
inline bool make_cache_map(CUtensorMap* map, const void* cache, int ces, int head_size,
                           int num_kv_heads, int nw, int64_t block_stride, int64_t token_stride,
                           int64_t head_stride) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) return false;
  const cuuint64_t dims[4] = {(cuuint64_t)head_size, (cuuint64_t)kBS, (cuuint64_t)num_kv_heads,
                              (cuuint64_t)1 << 31};
  const cuuint64_t strides[3] = {(cuuint64_t)token_stride * ces, (cuuint64_t)head_stride * ces,
                                 (cuuint64_t)block_stride * ces};
  const cuuint32_t box[4] = {(cuuint32_t)(128 / ces), (cuuint32_t)kBS, (cuuint32_t)nw, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult rc = fn(map, ces == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_UINT16,
                         4, const_cast<void*>(cache), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return rc == CUDA_SUCCESS;
}

template <int ST, bool FP8, bool E5, int D, int NW, int STAGES, int BPS, bool XP, int MINB>
int launch_nhd_t(const NhdParams& p_in, void* workspace, int64_t workspace_bytes, int num_sms,
                 cudaStream_t stream) {
  constexpr int CES = FP8 ? 1 : 2;
  constexpr int SMEM = STAGES * BPS * 2 * (D * CES * kBS * NW) + 1024;
  auto kern = paged_attention_nhd_kernel<ST, FP8, E5, D, NW, STAGES, BPS, XP, MINB>;
  static_assert(MINB * (SMEM + 2048) <= 232448, "the ring of MINB CTAs fits one SM");
  static bool attr_set[64] = {false};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
    return set_error(B200PA_ERR_CUDA, "cudaGetDevice failed or device ordinal out of range");
  if (!attr_set[dev]) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess)
      return set_error(B200PA_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    attr_set[dev] = true;
  }
  const int ctas = num_sms * MINB;
  if (ctas * (int)sizeof(int) > kCounterBytes)
    return set_error(B200PA_ERR_UNSUPPORTED, "nhd kernel: %d CTAs exceed the workspace header", ctas);
  const int64_t need = kCounterBytes + (int64_t)2 * ctas * NW * slot_floats(D) * (int64_t)sizeof(float);
  if (workspace_bytes < need)
    return set_error(B200PA_ERR_INVALID, "workspace too small: %lld bytes, need %lld",
                     (long long)workspace_bytes, (long long)need);
  CUtensorMap map_k, map_v;
  memset(&map_k, 0, sizeof(map_k));
  memset(&map_v, 0, sizeof(map_v));
  if constexpr (!XP) {
    if (!make_cache_map(&map_k, p_in.k_cache, CES, D, p_in.num_kv_heads, NW, p_in.kv_block_stride,
                        p_in.kv_token_stride, p_in.kv_head_stride) ||
        !make_cache_map(&map_v, p_in.v_cache, CES, D, p_in.num_kv_heads, NW, p_in.kv_block_stride,
                        p_in.kv_token_stride, p_in.kv_head_stride))
      return -1;     // the caller falls back to the stream kernel's row copies
  }
  NhdParams p = p_in;
  p.counters = reinterpret_cast<int*>(workspace);
  p.ws = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kCounterBytes);
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
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, map_k, map_v, p);
  if (e != cudaSuccess)
    return set_error(B200PA_ERR_CUDA, "nhd kernel launch: %s", cudaGetErrorString(e));
  count_launch();
  return check_launch("paged_attention (nhd)");
}
