// int_peak.cu -- micro-benchmark of the sustained integer-pipe rates of one GPU:
// the INT roofline denominators of the motion-search kernels (SURVEY.md 8d asks for
// a measured INT32 peak; MEASURED_PEAKS.json only holds HBM and bf16 tensor peaks).
// Each kernel runs dependency-free streams of one instruction on every SM and
// reports giga lane-operations per second.  Build: nvcc -gencode
// arch=compute_100a,code=sm_100a -O3 -o int_peak int_peak.cu ; prints one JSON line.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

constexpr int ITERS = 4096, ILP = 8, THREADS = 256;

#define OPK(NAME, BODY)                                                               \
  __global__ void __launch_bounds__(THREADS) NAME(uint32_t *out, uint32_t seed) {      \
    uint32_t a[ILP], b = seed + threadIdx.x, c = seed * 3 + 1;                         \
    _Pragma("unroll") for (int i = 0; i < ILP; i++) a[i] = seed + i * 977 + threadIdx.x; \
    for (int it = 0; it < ITERS; it++) {                                               \
      _Pragma("unroll") for (int i = 0; i < ILP; i++) { BODY; }                       \
    }                                                                                  \
    uint32_t s = 0;                                                                    \
    _Pragma("unroll") for (int i = 0; i < ILP; i++) s ^= a[i];                        \
    if (s == 0x12345678u) out[threadIdx.x] = s;                                        \
  }

OPK(k_iadd3, asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b)))
OPK(k_lop3, asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c)))
OPK(k_imad, asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c)))
OPK(k_vabsdiff4, asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c)))
OPK(k_dp4a, asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c)))
OPK(k_shf, asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c)))
OPK(k_prmt, asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c)))
OPK(k_min, asm volatile("min.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b)))
OPK(k_vabsdiff, asm volatile("vabsdiff.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c)))

__global__ void __launch_bounds__(THREADS) k_lds(uint32_t *out, uint32_t seed) {
  __shared__ uint32_t sm[THREADS * 2];
  sm[threadIdx.x] = seed;
  sm[threadIdx.x + THREADS] = seed;
  __syncthreads();
  uint32_t a[ILP];
  uint32_t idx = threadIdx.x;
#pragma unroll
  for (int i = 0; i < ILP; i++) a[i] = 0;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) a[i] += ((volatile uint32_t *)sm)[(idx + i * 32) & (2 * THREADS - 1)];
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s ^= a[i];
  if (s == 0x12345678u) out[threadIdx.x] = s;
}

template <typename K>
double run(K kern, int sms, uint32_t *d) {
  const int grid = sms * 8;
  kern<<<grid, THREADS>>>(d, 1);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    kern<<<grid, THREADS>>>(d, rep + 2);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double gops = (double)grid * THREADS * ITERS * ILP / (ms * 1e-3) / 1e9;
    if (gops > best) best = gops;
  }
  return best;
}

int main() {
  int sms = 0, clk = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  uint32_t *d;
  cudaMalloc(&d, 4096);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz\": %d, \"unit\": \"G lane-ops/s\"", p.name, sms, clk);
  printf(", \"iadd\": %.1f", run(k_iadd3, sms, d));
  printf(", \"lop3\": %.1f", run(k_lop3, sms, d));
  printf(", \"imad\": %.1f", run(k_imad, sms, d));
  printf(", \"vabsdiff4_acc\": %.1f", run(k_vabsdiff4, sms, d));
  printf(", \"vabsdiff_acc\": %.1f", run(k_vabsdiff, sms, d));
  printf(", \"dp4a\": %.1f", run(k_dp4a, sms, d));
  printf(", \"shf\": %.1f", run(k_shf, sms, d));
  printf(", \"prmt\": %.1f", run(k_prmt, sms, d));
  printf(", \"min_u32\": %.1f", run(k_min, sms, d));
  printf(", \"lds32\": %.1f", run(k_lds, sms, d));
  printf("}\n");
  cudaError_t e = cudaDeviceSynchronize();
  return e == cudaSuccess ? 0 : 1;
}
