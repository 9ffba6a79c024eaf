// diag.cu -- measured integer-pipe rates of the device a context runs on: the INT roofline denominators of the
// motion-search kernels (SURVEY 8d asks for a MEASURED INT32 peak; MEASURED_PEAKS.json holds only HBM and bf16 tensor
// peaks).  bench.py calls r1c_int_peak() in the same run as the kernels it rates.  Each kernel runs dependency-free
// streams (8 per thread) of one instruction on every SM; the result is giga lane-operations per second.
#include "common.cuh"

namespace {

constexpr int ITERS = 4096, ILP = 8, DT = 256;

#define OPK(NAME, BODY)                                                                   \
  __global__ void __launch_bounds__(DT) NAME(uint32_t *out, uint32_t seed) {              \
    uint32_t a[ILP], b = seed + threadIdx.x, c = seed * 3 + 1;                             \
    _Pragma("unroll") for (int i = 0; i < ILP; i++) a[i] = seed + i * 977 + threadIdx.x;   \
    for (int it = 0; it < ITERS; it++) {                                                   \
      _Pragma("unroll") for (int i = 0; i < ILP; i++) { BODY; }                           \
    }                                                                                      \
    uint32_t s = 0;                                                                        \
    _Pragma("unroll") for (int i = 0; i < ILP; i++) s ^= a[i];                            \
    if (s == 0x12345678u) out[threadIdx.x] = s;                                            \
  }

OPK(k_iadd3, asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b)))
OPK(k_vabsdiff4, asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c)))
OPK(k_dp4a, asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c)))
OPK(k_dp2a, asm volatile("dp2a.lo.u32.s32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c)))
OPK(k_imad, asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c)))
OPK(k_prmt, asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c)))

template <typename K>
int measure(r1c_ctx *ctx, K kern, uint32_t *d, double *gops) {
  const int grid = ctx->sm_count * 8;
  cudaEvent_t e0, e1;
  R1C_CUDA(ctx, cudaEventCreate(&e0));
  R1C_CUDA(ctx, cudaEventCreate(&e1));
  kern<<<grid, DT, 0, ctx->stream>>>(d, 1);
  double best = 0;
  for (int rep = 0; rep < 5; rep++) {
    R1C_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
    kern<<<grid, DT, 0, ctx->stream>>>(d, rep + 2);
    R1C_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
    R1C_CUDA(ctx, cudaEventSynchronize(e1));
    float ms = 0;
    R1C_CUDA(ctx, cudaEventElapsedTime(&ms, e0, e1));
    const double g = (double)grid * DT * ITERS * ILP / (ms * 1e-3) / 1e9;
    if (g > best) best = g;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  ctx->launches += 6;
  *gops = best;
  return R1C_OK;
}

}  // namespace

extern "C" int r1c_int_peak(r1c_ctx *ctx, int which, double *gops) {
  if (!ctx || !gops) return R1C_ERR_ARG;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  void *d;
  int rc = r1c_scratch_dev(ctx, 0, 4096, &d);
  if (rc) return rc;
  switch (which) {
    case R1C_PEAK_IADD: return measure(ctx, k_iadd3, (uint32_t *)d, gops);
    case R1C_PEAK_VABSDIFF4: return measure(ctx, k_vabsdiff4, (uint32_t *)d, gops);
    case R1C_PEAK_DP4A: return measure(ctx, k_dp4a, (uint32_t *)d, gops);
    case R1C_PEAK_DP2A: return measure(ctx, k_dp2a, (uint32_t *)d, gops);
    case R1C_PEAK_IMAD: return measure(ctx, k_imad, (uint32_t *)d, gops);
    case R1C_PEAK_PRMT: return measure(ctx, k_prmt, (uint32_t *)d, gops);
    default: return r1c_set_err(ctx, R1C_ERR_ARG, "r1c_int_peak: unknown instruction class %d", which);
  }
}
