/*
 * TEST INFRASTRUCTURE ONLY -- CPU oracle for the paged-decode hot path.
 *
 * Plain-C restatement of the reference's algorithm (Abandooon/vllm-main-modify,
 * paths relative to /root/reference).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may call into this file.  The
 * product (vllm_main_modify_b200/) never links, imports or executes it.
 *
 * Parity pin (see DESIGN.md "Oracle"): checked in tests/test_oracle.py against
 *   - golden FP8 bytes produced by the CUDA toolkit's own host implementation of
 *     __nv_cvt_float_to_fp8 / __nv_cvt_fp8_to_halfraw (cuda_fp8.hpp 12.9, the
 *     third-party dependency the reference calls at
 *     csrc/quantization/fp8/nvidia/quant_utils.cuh:311,368,475,487,500),
 *   - golden attention outputs produced by the reference's own test oracle
 *     `ref_single_query_cached_kv_attention`
 *     (tests/kernels/attention/test_attention.py:50-118) and by the reference's
 *     own CPU kernels csrc/cpu/attention.cpp / csrc/cpu/cache.cpp compiled
 *     unmodified into oracle/_ref (tests/golden/make_golden.py),
 *   - oracle/_ref itself, live, whenever it is present.
 *
 * What is restated:
 *   ora_f32_to_fp8 / ora_fp8_to_f32      quant_utils.cuh:300-512 (scaled_vec_conversion)
 *   ora_reshape_and_cache                cache_kernels.cu:212-264, 416-442
 *   ora_reshape_and_cache_flash          cache_kernels.cu:280-353, 459-497
 *   ora_convert_fp8                      cache_kernels.cu:552-623
 *   ora_paged_attention                  attention/attention_kernels.cuh:85-491 (v1/v2 body)
 *                                        and :560-664 (v2 reduce)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#ifdef _OPENMP
#include <omp.h>
#endif

enum { ORA_F32 = 0, ORA_F16 = 1, ORA_BF16 = 2 };           /* scalar_t            */
enum { ORA_KV_AUTO = 0, ORA_KV_E4M3 = 1, ORA_KV_E5M2 = 2 }; /* Fp8KVCacheDataType  */

/* ------------------------------------------------------------------ scalars */

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

float ora_bf16_to_f32(uint16_t h) { return u2f((uint32_t)h << 16); }

/* __float2bfloat16 (round-to-nearest-even, NaN -> 0x7FFF). */
uint16_t ora_f32_to_bf16(float f) {
  uint32_t u = f2u(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fff;
  uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return (uint16_t)(u >> 16);
}

float ora_f16_to_f32(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  int e = (h >> 10) & 0x1f;
  uint32_t m = h & 0x3ffu;
  if (e == 0) {
    if (m == 0) return u2f(sign);
    float v = ldexpf((float)m, -24);
    return sign ? -v : v;
  }
  if (e == 31) return u2f(sign | 0x7f800000u | (m << 13));
  return u2f(sign | ((uint32_t)(e + 112) << 23) | (m << 13));
}

/* cvt.rn.f16.f32 */
uint16_t ora_f32_to_f16(float f) {
  uint32_t u = f2u(f);
  uint16_t sign = (uint16_t)((u >> 16) & 0x8000u);
  uint32_t a = u & 0x7fffffffu;
  if (a > 0x7f800000u) return (uint16_t)(sign | 0x7fffu);
  if (a == 0x7f800000u) return (uint16_t)(sign | 0x7c00u);
  float af = u2f(a);
  if (af >= 65520.0f) return (uint16_t)(sign | 0x7c00u);
  if (af < 6.103515625e-05f) { /* subnormal half: step 2^-24 */
    float q = af * 16777216.0f; /* exact scaling by 2^24 */
    uint32_t r = (uint32_t)nearbyintf(q);
    return (uint16_t)(sign | r);
  }
  int e;
  float m = frexpf(af, &e);                /* af = m * 2^e, m in [0.5,1) */
  float q = nearbyintf(m * 2048.0f);       /* 11 significant bits */
  if (q == 2048.0f) { q = 1024.0f; e += 1; }
  uint32_t he = (uint32_t)(e - 1 + 15);
  if (he >= 31) return (uint16_t)(sign | 0x7c00u);
  return (uint16_t)(sign | (he << 10) | ((uint32_t)q - 1024u));
}

/* Value of an fp8 code.  The reference goes fp8 -> half (exact for both
 * formats) -> float (quant_utils.cuh:311-312, 368-371, 423-427). */
float ora_fp8_to_f32(uint8_t b, int kind) {
  int sign = b >> 7;
  float v;
  if (kind == ORA_KV_E5M2) {
    return ora_f16_to_f32((uint16_t)((uint16_t)b << 8));
  }
  int e = (b >> 3) & 0xf;
  int m = b & 7;
  if (e == 15 && m == 7) return u2f(sign ? 0xffffffffu : 0x7fffffffu);
  if (e == 0) v = ldexpf((float)m, -9);
  else v = ldexpf((float)(8 + m), e - 10);
  return sign ? -v : v;
}

/* __nv_cvt_float_to_fp8(x, __NV_SATFINITE, kind): round-to-nearest-even,
 * saturate to +-maxnorm (448 / 57344), NaN -> 0x7F (canonical, sign dropped:
 * the toolkit canonicalises NaN to 0x7FFFFFFF before converting).
 * Written from the format definition, not from the toolkit's bit routine. */
uint8_t ora_f32_to_fp8(float x, int kind) {
  const int mbits = (kind == ORA_KV_E5M2) ? 2 : 3;
  const int bias = (kind == ORA_KV_E5M2) ? 15 : 7;
  const double maxnorm = (kind == ORA_KV_E5M2) ? 57344.0 : 448.0;
  const uint8_t maxcode = (kind == ORA_KV_E5M2) ? 0x7B : 0x7E;
  const int emin = 1 - bias; /* exponent of the smallest normal */
  uint32_t u = f2u(x);
  uint8_t sign = (uint8_t)((u >> 31) << 7);
  if ((u & 0x7fffffffu) > 0x7f800000u) {
    return (uint8_t)0x7F;
  }
  double a = fabs((double)x);
  if (a == 0.0) return sign;
  /* quantum of the binade holding a (subnormals share the emin quantum) */
  int e;
  (void)frexp(a, &e); /* a = f * 2^e, f in [0.5, 1)  -> exponent of a is e-1 */
  int ex = e - 1;
  if (ex < emin) ex = emin;
  double quantum = ldexp(1.0, ex - mbits);
  double q = nearbyint(a / quantum); /* ties-to-even in the default mode */
  double r = q * quantum;
  if (r > maxnorm || isinf(a)) return (uint8_t)(maxcode | sign);
  if (r == 0.0) return sign;
  (void)frexp(r, &e);
  ex = e - 1;
  uint8_t code;
  if (ex < emin) {
    code = (uint8_t)(r / ldexp(1.0, emin - mbits));
  } else {
    uint32_t mant = (uint32_t)(r / ldexp(1.0, ex - mbits)) - (1u << mbits);
    code = (uint8_t)((((uint32_t)(ex + bias)) << mbits) | mant);
  }
  return (uint8_t)(code | sign);
}

/* Load element i of a scalar_t array as float. */
static inline float ld_scalar(const void* p, int64_t i, int st) {
  if (st == ORA_F32) return ((const float*)p)[i];
  if (st == ORA_F16) return ora_f16_to_f32(((const uint16_t*)p)[i]);
  return ora_bf16_to_f32(((const uint16_t*)p)[i]);
}
/* from_float(): round a float to scalar_t and return the rounded value. */
static inline float round_scalar(float v, int st) {
  if (st == ORA_F32) return v;
  if (st == ORA_F16) return ora_f16_to_f32(ora_f32_to_f16(v));
  return ora_bf16_to_f32(ora_f32_to_bf16(v));
}
static inline void st_scalar(void* p, int64_t i, float v, int st) {
  if (st == ORA_F32) ((float*)p)[i] = v;
  else if (st == ORA_F16) ((uint16_t*)p)[i] = ora_f32_to_f16(v);
  else ((uint16_t*)p)[i] = ora_f32_to_bf16(v);
}
static inline int scalar_size(int st) { return st == ORA_F32 ? 4 : 2; }

/* scaled_convert<cache_t=uint8_t, scalar_t>(x, scale): fp8(float(x) / scale),
 * true IEEE fp32 division (quant_utils.cuh:469-502). */
static inline uint8_t quant_scalar(float x, float scale, int kind) {
  volatile float q = x / scale;
  return ora_f32_to_fp8(q, kind);
}
/* scaled_convert<scalar_t, uint8_t>(b, scale): scalar_t(float(half(b)) * scale)
 * (quant_utils.cuh:306-313 half, :361-372 bf16, :418-428 float). */
static inline float dequant_scalar(uint8_t b, float scale, int kind, int st) {
  volatile float v = ora_fp8_to_f32(b, kind) * scale;
  return round_scalar(v, st);
}

/* --------------------------------------------------------- cache write ops */

/* reshape_and_cache (cache_kernels.cu:212-264; host wrapper :416-442).
 * key/value: [num_tokens, num_heads, head_size] with row strides in elements.
 * key_cache: [num_blocks, num_heads, head_size/x, block_size, x]
 * value_cache: [num_blocks, num_heads, head_size, block_size]
 * The cache is indexed densely exactly like the reference kernel (:244-251). */
int ora_reshape_and_cache(const void* key, const void* value, void* key_cache,
                          void* value_cache, const int64_t* slot_mapping,
                          int num_tokens, int64_t key_stride, int64_t value_stride,
                          int num_heads, int head_size, int block_size, int x,
                          int scalar_type, int kv_dtype, float k_scale,
                          float v_scale) {
  const int n = num_heads * head_size;
  for (int64_t t = 0; t < num_tokens; ++t) {
    const int64_t slot = slot_mapping[t];
    if (slot < 0) continue; /* padding token (:226) */
    const int64_t block_idx = slot / block_size;
    const int64_t block_off = slot % block_size;
    for (int i = 0; i < n; ++i) {
      const int head = i / head_size, ho = i % head_size;
      const int xi = ho / x, xo = ho % x;
      const int64_t kdst =
          block_idx * num_heads * (head_size / x) * block_size * x +
          (int64_t)head * (head_size / x) * block_size * x +
          (int64_t)xi * block_size * x + block_off * x + xo;
      const int64_t vdst = block_idx * num_heads * head_size * block_size +
                           (int64_t)head * head_size * block_size +
                           (int64_t)ho * block_size + block_off;
      if (kv_dtype == ORA_KV_AUTO) {
        const int es = scalar_size(scalar_type);
        memcpy((char*)key_cache + kdst * es,
               (const char*)key + (t * key_stride + i) * es, es);
        memcpy((char*)value_cache + vdst * es,
               (const char*)value + (t * value_stride + i) * es, es);
      } else {
        ((uint8_t*)key_cache)[kdst] = quant_scalar(
            ld_scalar(key, t * key_stride + i, scalar_type), k_scale, kv_dtype);
        ((uint8_t*)value_cache)[vdst] = quant_scalar(
            ld_scalar(value, t * value_stride + i, scalar_type), v_scale, kv_dtype);
      }
    }
  }
  return 0;
}

/* reshape_and_cache_flash (cache_kernels.cu:280-353; wrapper :459-497).
 * Cache is addressed purely by strides (elements): NHD has
 * head_stride == head_size, HND has page_stride == head_size. */
int ora_reshape_and_cache_flash(const void* key, const void* value,
                                void* key_cache, void* value_cache,
                                const int64_t* slot_mapping, int num_tokens,
                                int64_t key_stride, int64_t value_stride,
                                int64_t block_stride, int64_t page_stride,
                                int64_t head_stride, int num_heads,
                                int head_size, int block_size, int scalar_type,
                                int kv_dtype, float k_scale, float v_scale) {
  for (int64_t t = 0; t < num_tokens; ++t) {
    const int64_t slot = slot_mapping[t];
    if (slot < 0) continue;
    const int64_t block_idx = slot / block_size;
    const int64_t block_off = slot % block_size;
    for (int h = 0; h < num_heads; ++h) {
      for (int d = 0; d < head_size; ++d) {
        const int64_t src = (int64_t)h * head_size + d;
        const int64_t dst = block_idx * block_stride + block_off * page_stride +
                            (int64_t)h * head_stride + d;
        if (kv_dtype == ORA_KV_AUTO) {
          const int es = scalar_size(scalar_type);
          memcpy((char*)key_cache + dst * es,
                 (const char*)key + (t * key_stride + src) * es, es);
          memcpy((char*)value_cache + dst * es,
                 (const char*)value + (t * value_stride + src) * es, es);
        } else {
          ((uint8_t*)key_cache)[dst] = quant_scalar(
              ld_scalar(key, t * key_stride + src, scalar_type), k_scale, kv_dtype);
          ((uint8_t*)value_cache)[dst] = quant_scalar(
              ld_scalar(value, t * value_stride + src, scalar_type), v_scale,
              kv_dtype);
        }
      }
    }
  }
  return 0;
}

/* convert_fp8 (cache_kernels.cu:552-623): elementwise over n elements.
 * to_fp8 != 0: src scalar_t -> dst uint8;  else src uint8 -> dst scalar_t. */
int ora_convert_fp8(void* dst, const void* src, int64_t n, float scale,
                    int scalar_type, int kv_dtype, int to_fp8) {
  if (kv_dtype != ORA_KV_E4M3 && kv_dtype != ORA_KV_E5M2) return 1;
  for (int64_t i = 0; i < n; ++i) {
    if (to_fp8) {
      ((uint8_t*)dst)[i] = quant_scalar(ld_scalar(src, i, scalar_type), scale, kv_dtype);
    } else {
      volatile float v = ora_fp8_to_f32(((const uint8_t*)src)[i], kv_dtype) * scale;
      st_scalar(dst, i, v, scalar_type);
    }
  }
  return 0;
}

/* -------------------------------------------------------- paged attention */

typedef struct {
  const void* key_cache;
  const void* value_cache;
  int64_t kv_block_stride, kv_head_stride;
  int head_size, block_size, x;
  int scalar_type, kv_dtype;
  float k_scale, v_scale;
  int faithful;
} kv_view;

/* K element d of token slot `off` in physical block `blk`, kv head `h`
 * (address math of attention_kernels.cuh:268-281). */
static inline double load_k(const kv_view* c, int64_t blk, int h, int off, int d) {
  const int64_t idx = blk * c->kv_block_stride + (int64_t)h * c->kv_head_stride +
                      (int64_t)(d / c->x) * c->block_size * c->x +
                      (int64_t)off * c->x + (d % c->x);
  if (c->kv_dtype == ORA_KV_AUTO) return ld_scalar(c->key_cache, idx, c->scalar_type);
  const uint8_t b = ((const uint8_t*)c->key_cache)[idx];
  if (c->faithful) return dequant_scalar(b, c->k_scale, c->kv_dtype, c->scalar_type);
  return (double)ora_fp8_to_f32(b, c->kv_dtype) * (double)c->k_scale;
}
/* V element d of token slot `off` (attention_kernels.cuh:397-413). */
static inline double load_v(const kv_view* c, int64_t blk, int h, int off, int d) {
  const int64_t idx = blk * c->kv_block_stride + (int64_t)h * c->kv_head_stride +
                      (int64_t)d * c->block_size + off;
  if (c->kv_dtype == ORA_KV_AUTO) return ld_scalar(c->value_cache, idx, c->scalar_type);
  const uint8_t b = ((const uint8_t*)c->value_cache)[idx];
  if (c->faithful) return dequant_scalar(b, c->v_scale, c->kv_dtype, c->scalar_type);
  return (double)ora_fp8_to_f32(b, c->kv_dtype) * (double)c->v_scale;
}

/*
 * One (seq, head, token range) pass: attention_kernels.cuh:85-491.
 * Writes `acc[head_size]` (normalised partial output), *max_logit, *exp_sum.
 *   faithful=1: fp32 logits, expf, 1/(sum+1e-6f), probabilities rounded to
 *               scalar_t before P.V (:393-395), fp32 accumulators.
 *   faithful=0: the same formula in double without intermediate rounding.
 */
static void attend_range(const kv_view* c, const void* query, int64_t q_off,
                         const int32_t* block_table, int seq_len, int tok_begin,
                         int tok_end, int kv_head, float scale, float alibi_slope,
                         int sparse, int bs_block_offset, int q_bs_block_id,
                         int bs_local_blocks, int bs_vert_stride,
                         int bs_block_size, double* logits, double* acc,
                         float* max_logit, float* exp_sum) {
  const int D = c->head_size, BS = c->block_size;
  const int n = tok_end - tok_begin;
  double qmax = -FLT_MAX;
  for (int i = 0; i < n; ++i) {
    const int tok = tok_begin + i;
    const int block_idx = tok / BS;
    if (sparse) {
      const int kb = block_idx * BS / bs_block_size;
      const int is_remote = ((kb + bs_block_offset) % bs_vert_stride == 0);
      const int is_local = (kb > q_bs_block_id - bs_local_blocks);
      if (!is_remote && !is_local) { logits[i] = -FLT_MAX; continue; } /* :236-249 */
    }
    const int64_t blk = block_table[block_idx];
    double dot = 0.0;
    for (int d = 0; d < D; ++d)
      dot += (double)ld_scalar(query, q_off + d, c->scalar_type) *
             load_k(c, blk, kv_head, tok % BS, d);
    double qk;
    if (c->faithful) {
      float f = scale * (float)dot;                                   /* :289 */
      if (alibi_slope != 0.f) f += alibi_slope * (float)(tok - seq_len + 1); /* :292 */
      qk = f;
    } else {
      qk = (double)scale * dot;
      if (alibi_slope != 0.f) qk += (double)alibi_slope * (double)(tok - seq_len + 1);
    }
    logits[i] = qk;
    if (qk > qmax) qmax = qk;
  }
  double sum = 0.0;
  for (int i = 0; i < n; ++i) {
    double e = c->faithful ? (double)expf((float)logits[i] - (float)qmax)
                           : exp(logits[i] - qmax);
    logits[i] = e;
    sum += e;
  }
  if (c->faithful) sum = (double)(float)sum;
  const double inv = c->faithful ? (double)(1.0f / ((float)sum + 1e-6f))
                                 : 1.0 / (sum + 1e-6);               /* :337 */
  for (int d = 0; d < D; ++d) acc[d] = 0.0;
  for (int i = 0; i < n; ++i) {
    const int tok = tok_begin + i;
    const int block_idx = tok / BS;
    if (sparse) {
      const int vb = block_idx * BS / bs_block_size;
      if (!((vb + bs_block_offset) % bs_vert_stride == 0) &&
          !(vb > q_bs_block_id - bs_local_blocks))
        continue;                                                     /* :382-387 */
    }
    double p = logits[i] * inv;
    if (c->faithful) p = round_scalar((float)((float)logits[i] * (float)inv), c->scalar_type);
    const int64_t blk = block_table[block_idx];
    for (int d = 0; d < D; ++d) acc[d] += p * load_v(c, blk, kv_head, tok % BS, d);
  }
  *max_logit = (float)qmax;
  *exp_sum = (float)sum;
}

/*
 * paged_attention_v1 (partition_size == 0) / paged_attention_v2
 * (partition_size == 512).  Tensor layouts as vllm/_custom_ops.py:39-97.
 * out_f32 (optional, may be NULL) receives the unrounded result.
 * For v2, exp_sums / max_logits / tmp_out follow the reference layout
 * [num_seqs, num_heads, max_num_partitions(, head_size)].
 */
int ora_paged_attention(void* out, float* out_f32, const void* query,
                        int64_t q_stride, const void* key_cache,
                        const void* value_cache, int64_t kv_block_stride,
                        int64_t kv_head_stride, int num_seqs, int num_heads,
                        int num_kv_heads, int head_size, int block_size, int x,
                        float scale, const int32_t* block_tables,
                        int max_num_blocks_per_seq, const int32_t* seq_lens,
                        const float* alibi_slopes, int scalar_type, int kv_dtype,
                        float k_scale, float v_scale, int partition_size,
                        int faithful, float* exp_sums, float* max_logits,
                        void* tmp_out, int max_num_partitions, int tp_rank,
                        int bs_local_blocks, int bs_vert_stride,
                        int bs_block_size, int bs_head_sliding_step) {
  if (num_heads % num_kv_heads != 0) return 1;
  kv_view c = {key_cache, value_cache, kv_block_stride, kv_head_stride,
               head_size, block_size, x, scalar_type, kv_dtype,
               k_scale, v_scale, faithful};
  const int G = num_heads / num_kv_heads;
  const int sparse = bs_vert_stride > 1;                 /* paged_attention_v1.cu:175 */
  int max_len = 0;
  for (int s = 0; s < num_seqs; ++s) if (seq_lens[s] > max_len) max_len = seq_lens[s];
  int status = 0;

#pragma omp parallel for collapse(2) schedule(dynamic, 1)
  for (int s = 0; s < num_seqs; ++s) {
    for (int h = 0; h < num_heads; ++h) {
      const int seq_len = seq_lens[s];
      const int kvh = h / G;
      const float slope = alibi_slopes ? alibi_slopes[h] : 0.f;
      const int32_t* bt = block_tables + (int64_t)s * max_num_blocks_per_seq;
      int bs_off = 0, q_bs = 0;
      if (sparse) {
        q_bs = (seq_len - 1) / bs_block_size;
        if (bs_head_sliding_step >= 0)
          bs_off = (tp_rank * num_heads + h) * bs_head_sliding_step + 1;
        else
          bs_off = (tp_rank * num_kv_heads + kvh) * (-bs_head_sliding_step) + 1;
      }
      const int cap = partition_size > 0 ? partition_size : (seq_len > 0 ? seq_len : 1);
      double* logits = (double*)malloc(sizeof(double) * (size_t)cap);
      double* acc = (double*)malloc(sizeof(double) * (size_t)head_size);
      const int64_t q_off = (int64_t)s * q_stride + (int64_t)h * head_size;
      const int64_t o_off = ((int64_t)s * num_heads + h) * head_size;
      if (partition_size == 0) {
        float m, l;
        attend_range(&c, query, q_off, bt, seq_len, 0, seq_len, kvh, scale, slope,
                     sparse, bs_off, q_bs, bs_local_blocks, bs_vert_stride,
                     bs_block_size, logits, acc, &m, &l);
        for (int d = 0; d < head_size; ++d) {
          st_scalar(out, o_off + d, (float)acc[d], scalar_type);
          if (out_f32) out_f32[o_off + d] = (float)acc[d];
        }
      } else {
        const int np = (seq_len + partition_size - 1) / partition_size;
        const int64_t base = ((int64_t)s * num_heads + h) * max_num_partitions;
        for (int p = 0; p < np; ++p) {
          const int b = p * partition_size;
          const int e = b + partition_size < seq_len ? b + partition_size : seq_len;
          float m, l;
          attend_range(&c, query, q_off, bt, seq_len, b, e, kvh, scale, slope,
                       sparse, bs_off, q_bs, bs_local_blocks, bs_vert_stride,
                       bs_block_size, logits, acc, &m, &l);
          max_logits[base + p] = m;                                  /* :344-352 */
          exp_sums[base + p] = l;
          for (int d = 0; d < head_size; ++d)
            st_scalar(tmp_out, (base + p) * head_size + d, (float)acc[d], scalar_type);
        }
        /* reduce kernel :560-664 */
        if (!faithful) {
          /* exact mode: scratch is filled above, `out` is the unsplit result
           * (mathematically identical to the merge, without tmp_out rounding) */
          float m, l;
          double* lg = (double*)malloc(sizeof(double) * (size_t)(seq_len > 0 ? seq_len : 1));
          attend_range(&c, query, q_off, bt, seq_len, 0, seq_len, kvh, scale, slope,
                       sparse, bs_off, q_bs, bs_local_blocks, bs_vert_stride,
                       bs_block_size, lg, acc, &m, &l);
          free(lg);
          for (int d = 0; d < head_size; ++d) {
            st_scalar(out, o_off + d, (float)acc[d], scalar_type);
            if (out_f32) out_f32[o_off + d] = (float)acc[d];
          }
        } else if (np == 1) {
          for (int d = 0; d < head_size; ++d) {
            const float v = ld_scalar(tmp_out, base * head_size + d, scalar_type);
            st_scalar(out, o_off + d, v, scalar_type);                /* :577-589 */
            if (out_f32) out_f32[o_off + d] = v;
          }
        } else {
          float gmax = -FLT_MAX;
          for (int p = 0; p < np; ++p) if (max_logits[base + p] > gmax) gmax = max_logits[base + p];
          double gsum = 0.0;
          double* w = (double*)malloc(sizeof(double) * (size_t)np);
          for (int p = 0; p < np; ++p) {
            w[p] = faithful ? (double)(exp_sums[base + p] * expf(max_logits[base + p] - gmax))
                            : (double)exp_sums[base + p] * exp((double)max_logits[base + p] - (double)gmax);
            gsum += w[p];
          }
          const double inv = faithful ? (double)(1.0f / ((float)gsum + 1e-6f))
                                      : 1.0 / (gsum + 1e-6);          /* :647 */
          for (int d = 0; d < head_size; ++d) {
            double a = 0.0;
            for (int p = 0; p < np; ++p)
              a += (double)ld_scalar(tmp_out, (base + p) * head_size + d, scalar_type) *
                   w[p] * inv;                                         /* :656-663 */
            st_scalar(out, o_off + d, (float)a, scalar_type);
            if (out_f32) out_f32[o_off + d] = (float)a;
          }
          free(w);
        }
      }
      free(logits); free(acc);
    }
  }
  (void)max_len;
  return status;
}

int ora_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
