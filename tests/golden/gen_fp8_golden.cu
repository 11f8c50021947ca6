// Golden-vector generator (committed with the vectors it made; run by
// tests/golden/make_golden.py in the build container only).
//
// Calls the CUDA toolkit's OWN host implementations from cuda_fp8.hpp /
// cuda_fp16.hpp / cuda_bf16.hpp (CUDA 12.9.86) -- the third-party dependency the
// reference's quantise/dequantise primitives are written against
// (csrc/quantization/fp8/nvidia/quant_utils.cuh:311,368,475,487,500):
//   __nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3 | __NV_E5M2)
//   __nv_cvt_fp8_to_halfraw(b, kind)
//   __float2bfloat16 / __float2half_rn
// Output (binary, little endian) to stdout:
//   u8  e4m3_from_bf16[65536]   code of float(bf16 bits i)
//   u8  e5m2_from_bf16[65536]
//   u8  e4m3_from_f16[65536]    code of float(f16 bits i)
//   u8  e5m2_from_f16[65536]
//   u16 half_from_e4m3[256]     __nv_cvt_fp8_to_halfraw
//   u16 half_from_e5m2[256]
//   u16 bf16_from_f32[N], u16 f16_from_f32[N]  for the N probe floats below
//   u8  e4m3_scaled[3][65536]   code of float(bf16 bits i) / scale_j  (fp32 divide)
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

static float bf16_bits_to_float(uint16_t b) {
  uint32_t u = (uint32_t)b << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static float f16_bits_to_float(uint16_t b) {
  __half_raw r;
  r.x = b;
  return __half2float(__half(r));
}

int main() {
  std::vector<uint8_t> a(65536), b(65536), c(65536), d(65536);
  for (uint32_t i = 0; i < 65536; ++i) {
    float x = bf16_bits_to_float((uint16_t)i);
    a[i] = __nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3);
    b[i] = __nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E5M2);
    float y = f16_bits_to_float((uint16_t)i);
    c[i] = __nv_cvt_float_to_fp8(y, __NV_SATFINITE, __NV_E4M3);
    d[i] = __nv_cvt_float_to_fp8(y, __NV_SATFINITE, __NV_E5M2);
  }
  fwrite(a.data(), 1, 65536, stdout);
  fwrite(b.data(), 1, 65536, stdout);
  fwrite(c.data(), 1, 65536, stdout);
  fwrite(d.data(), 1, 65536, stdout);
  uint16_t h4[256], h5[256];
  for (int i = 0; i < 256; ++i) {
    h4[i] = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)i, __NV_E4M3).x;
    h5[i] = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)i, __NV_E5M2).x;
  }
  fwrite(h4, 2, 256, stdout);
  fwrite(h5, 2, 256, stdout);
  // probe floats: every bf16 pattern nudged by +-1 ulp-of-float and midpoints
  const int N = 65536 * 4;
  std::vector<uint16_t> pb(N), ph(N);
  for (uint32_t i = 0; i < 65536; ++i) {
    for (int j = 0; j < 4; ++j) {
      uint32_t u = ((uint32_t)i << 16) | (j == 0 ? 0x0000u : j == 1 ? 0x7fffu : j == 2 ? 0x8000u : 0x8001u);
      float f;
      memcpy(&f, &u, 4);
      pb[i * 4 + j] = __bfloat16_as_ushort(__float2bfloat16(f));
      ph[i * 4 + j] = __half_as_ushort(__float2half_rn(f));
    }
  }
  fwrite(pb.data(), 2, N, stdout);
  fwrite(ph.data(), 2, N, stdout);
  const float scales[3] = {0.5f, 0.0123f, 3.7f};
  for (int s = 0; s < 3; ++s) {
    for (uint32_t i = 0; i < 65536; ++i) {
      float x = bf16_bits_to_float((uint16_t)i) / scales[s];
      a[i] = __nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3);
    }
    fwrite(a.data(), 1, 65536, stdout);
  }
  return 0;
}
