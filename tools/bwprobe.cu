// HBM read-bandwidth probes for the roofline discussion in DESIGN.md (development tool).
//   A: coalesced 16-byte loads, grid-stride, read only
//   B: the decode kernel's access pattern without the math: per-warp rings filled by 1-D TMA
//      bulk copies of randomly permuted 2 KiB / 4 KiB runs, consumer only waits and releases
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bwprobe tools/bwprobe.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__global__ void read_plain(const uint4* __restrict__ src, size_t n, uint32_t* sink) {
  uint32_t acc = 0;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    uint4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
    acc ^= a.x ^ b.y ^ c.z ^ d.w;
  }
  for (; i < n; i += stride) acc ^= src[i].x;
  if (acc == 0x12345678u) *sink = acc;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int NW, int STAGES, int RUN>   // RUN bytes per bulk copy, 2 copies per stage
__global__ void __launch_bounds__(NW * 32) read_tma(const uint8_t* __restrict__ src, const int* __restrict__ perm,
                                                    int runs_per_warp, uint32_t* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[NW * STAGES];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t ring = smem_u32(smem) + warp * STAGES * 2 * RUN;
  const uint32_t bar0 = smem_u32(&bars[warp * STAGES]);
  if (lane == 0) {
    for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8 * s));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const long long wid = (long long)blockIdx.x * NW + warp;
  const int* my = perm + wid * runs_per_warp;
  const int nst = runs_per_warp / 2;
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  auto issue = [&](int i) {
    if (lane == 0) {
      const int s = i % STAGES;
      const uint32_t bar = bar0 + 8 * s, dst = ring + s * 2 * RUN;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(2 * RUN) : "memory");
      for (int g = 0; g < 2; ++g) {
        const uint8_t* p = src + (size_t)my[2 * i + g] * RUN;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                     ::"r"(dst + g * RUN), "l"(p), "r"(RUN), "r"(bar), "l"(pol) : "memory");
      }
    }
  };
  for (int i = 0; i < STAGES && i < nst; ++i) issue(i);
  uint32_t acc = 0;
  for (int i = 0; i < nst; ++i) {
    const int s = i % STAGES;
    const uint32_t bar = bar0 + 8 * s, par = (i / STAGES) & 1;
    asm volatile("{\n.reg .pred P1;\nW: mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra D;\nbra W;\nD:\n}\n" ::"r"(bar), "r"(par) : "memory");
    acc ^= *reinterpret_cast<volatile uint32_t*>(smem + (size_t)warp * STAGES * 2 * RUN + s * 2 * RUN + lane * 4);
    __syncwarp();
    if (i + STAGES < nst) issue(i + STAGES);
  }
  if (acc == 0x12345678u) *sink = acc;
}

template <typename F> float time_ms(F f, int iters) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  f(); f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int i = 0; i < iters; ++i) {
    CK(cudaEventRecord(a)); f(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b)); best = std::min(best, ms);
  }
  return best;
}

template <int NW, int STAGES, int RUN>
void run_tma(const uint8_t* buf, size_t bytes, int ctas_per_sm, int sms, uint32_t* sink, const char* name) {
  const size_t nruns = bytes / RUN;
  const int workers = sms * ctas_per_sm * NW;
  int rpw = (int)(nruns / workers); rpw &= ~1;
  std::vector<int> perm(nruns);
  for (size_t i = 0; i < nruns; ++i) perm[i] = (int)i;
  uint64_t st = 88172645463325252ull;
  for (size_t i = nruns - 1; i > 0; --i) { st ^= st << 13; st ^= st >> 7; st ^= st << 17; std::swap(perm[i], perm[st % (i + 1)]); }
  int* dperm; CK(cudaMalloc(&dperm, nruns * sizeof(int)));
  CK(cudaMemcpy(dperm, perm.data(), nruns * sizeof(int), cudaMemcpyHostToDevice));
  const int smem = NW * STAGES * 2 * RUN;
  auto k = read_tma<NW, STAGES, RUN>;
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  float ms = time_ms([&] { k<<<sms * ctas_per_sm, NW * 32, smem>>>(buf, dperm, rpw, sink); }, 10);
  CK(cudaGetLastError());
  const double got = (double)workers * rpw * RUN;
  printf("{\"probe\": \"%s\", \"run_bytes\": %d, \"warps_per_sm\": %d, \"stages\": %d, \"ring_kib_per_sm\": %d, \"ms\": %.4f, \"gbs\": %.1f}\n",
         name, RUN, ctas_per_sm * NW, STAGES, ctas_per_sm * smem / 1024, ms, got / ms / 1e6);
  CK(cudaFree(dperm));
}

// `bwprobe sustain [seconds]`: the 8-warp 2 KiB TMA ring probe (the decode kernel's traffic without
// its math) and the plain-load probe back to back for `seconds` each, one line per ~100 ms: does
// a pure read stream hold its rate for seconds, and at what power / SM clock (sample nvidia-smi
// beside it)?  The difference to the decode kernel's sustained leg is the cost of the math.
template <int NW, int STAGES, int RUN>
void sustain_tma(const uint8_t* buf, size_t bytes, int ctas_per_sm, int sms, uint32_t* sink, double seconds,
                 const char* name) {
  const size_t nruns = bytes / RUN;
  const int workers = sms * ctas_per_sm * NW;
  int rpw = (int)(nruns / workers); rpw &= ~1;
  std::vector<int> perm(nruns);
  for (size_t i = 0; i < nruns; ++i) perm[i] = (int)i;
  uint64_t st = 88172645463325252ull;
  for (size_t i = nruns - 1; i > 0; --i) { st ^= st << 13; st ^= st >> 7; st ^= st << 17; std::swap(perm[i], perm[st % (i + 1)]); }
  int* dperm; CK(cudaMalloc(&dperm, nruns * sizeof(int)));
  CK(cudaMemcpy(dperm, perm.data(), nruns * sizeof(int), cudaMemcpyHostToDevice));
  const int smem = NW * STAGES * 2 * RUN;
  auto k = read_tma<NW, STAGES, RUN>;
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const double got = (double)workers * rpw * RUN;
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  double elapsed = 0;
  while (elapsed < seconds * 1e3) {
    CK(cudaEventRecord(a));
    for (int i = 0; i < 80; ++i) k<<<sms * ctas_per_sm, NW * 32, smem>>>(buf, dperm, rpw, sink);
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    elapsed += ms;
    printf("{\"probe\": \"%s\", \"t_ms\": %.0f, \"gbs\": %.1f}\n", name, elapsed, 80 * got / ms / 1e6);
    fflush(stdout);
  }
  CK(cudaFree(dperm));
}

int main(int argc, char** argv) {
  const size_t bytes = (size_t)8 << 30;
  uint8_t* buf; CK(cudaMalloc(&buf, bytes)); CK(cudaMemset(buf, 1, bytes));
  uint32_t* sink; CK(cudaMalloc(&sink, 4));
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  if (argc > 1 && !strcmp(argv[1], "sustain")) {
    const double seconds = argc > 2 ? atof(argv[2]) : 4.0;
    sustain_tma<8, 6, 2048>(buf, bytes, 1, sms, sink, seconds, "sustain_tma_2k_8w_6st");
    double elapsed = 0;
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    while (elapsed < seconds * 1e3) {
      CK(cudaEventRecord(a));
      for (int i = 0; i < 80; ++i) read_plain<<<sms * 8, 256>>>((const uint4*)buf, bytes / 16, sink);
      CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
      float ms; CK(cudaEventElapsedTime(&ms, a, b));
      elapsed += ms;
      printf("{\"probe\": \"sustain_plain_ld128\", \"t_ms\": %.0f, \"gbs\": %.1f}\n", elapsed, 80.0 * bytes / ms / 1e6);
      fflush(stdout);
    }
    return 0;
  }
  for (int cps : {4, 8, 16}) {
    float ms = time_ms([&] { read_plain<<<sms * cps, 256>>>((const uint4*)buf, bytes / 16, sink); }, 10);
    printf("{\"probe\": \"plain_ld128\", \"ctas_per_sm\": %d, \"ms\": %.4f, \"gbs\": %.1f}\n", cps, ms, bytes / ms / 1e6);
  }
  run_tma<4, 6, 2048>(buf, bytes, 2, sms, sink, "tma_2k_8w_6st");
  run_tma<4, 3, 4096>(buf, bytes, 2, sms, sink, "tma_4k_8w_3st");
  run_tma<4, 4, 4096>(buf, bytes, 1, sms, sink, "tma_4k_4w_4st");
  run_tma<4, 6, 4096>(buf, bytes, 1, sms, sink, "tma_4k_4w_6st");
  run_tma<4, 2, 2048>(buf, bytes, 4, sms, sink, "tma_2k_16w_2st");
  run_tma<4, 12, 2048>(buf, bytes, 1, sms, sink, "tma_2k_4w_12st");
  run_tma<2, 12, 4096>(buf, bytes, 1, sms, sink, "tma_4k_2w_12st");
  run_tma<4, 2, 16384>(buf, bytes, 1, sms, sink, "tma_16k_4w_2st");
  return 0;
}
