// tma_probe.cu -- minimal 2-D TMA box load (debugging aid for mc_tma.cu): nvcc -arch=sm_100a -o build/tools/tma_probe tools/tma_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
struct Args {
  alignas(64) CUtensorMap map[2];
  uint16_t *out;
  const CUtensorMap *gmap;
  const uint16_t *src;
  int x, y, variant, bytes, bw;
};
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(const __grid_constant__ Args a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t *bar = (uint64_t *)(smem + 8192);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(bar)), "r"(32));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int lane = threadIdx.x;
  if (lane < 32) {
    if (a.variant == 0) {
      if (lane == 0) asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(s32(bar)), "r"(2 * 32 * 15) : "memory");
      if (lane < 2)
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                         s32(smem + lane * 512)),
                     "l"((unsigned long long)&a.map[lane & 1]), "r"(a.x + lane), "r"(a.y), "r"(s32(bar))
                     : "memory");
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
    } else if (a.variant == 3) {
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
    } else if (a.variant == 4) {
      if (lane == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(480) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(smem)),
                     "l"(__cvta_generic_to_global(a.src)), "r"(480), "r"(s32(bar)) : "memory");
      } else {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
      }
    } else if (a.variant == 5) {
      if (lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"((unsigned long long)&a.map[0]) : "memory");
      }
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
    } else {
      if (lane == 0) {
        const CUtensorMap *mp = a.variant == 2 ? a.gmap : &a.map[0];
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(a.bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                         s32(smem)),
                     "l"((unsigned long long)mp), "r"(a.x), "r"(a.y), "r"(s32(bar))
                     : "memory");
      } else {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
      }
    }
  }
  asm volatile(
      "{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(s32(bar)), "r"(0)
      : "memory");
  for (int i = threadIdx.x; i < 16 * 15; i += blockDim.x) a.out[i] = ((uint16_t *)smem)[i];
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char **argv) {
  const int variant = argc > 1 ? atoi(argv[1]) : 0;
  const int W = 4096, H = 256;
  uint16_t *h = (uint16_t *)malloc(W * H * 2), *d, *out;
  for (int i = 0; i < W * H; i++) h[i] = (uint16_t)((i / W) * 100 + (i % W) % 100);
  cudaMalloc(&d, W * H * 2);
  cudaMalloc(&out, 4096);
  cudaMemcpy(d, h, W * H * 2, cudaMemcpyHostToDevice);
  void *fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  printf("entry point: %d %d %p\n", (int)e, (int)q, fn);
  Args a;
  const cuuint64_t dims[2] = {W, H}, strides[1] = {W * 2};
  const int bw = argc > 2 ? atoi(argv[2]) : 16, bh = argc > 3 ? atoi(argv[3]) : 15, promo = argc > 4 ? atoi(argv[4]) : 2, swz = argc > 5 ? atoi(argv[5]) : 0;
  const cuuint32_t box[2] = {(cuuint32_t)bw, (cuuint32_t)bh}, es[2] = {1, 1};
  printf("box %d x %d promo %d\n", bw, bh, promo);
  for (int k = 0; k < 2; k++) {
    CUresult r = ((EncodeTiledFn)fn)(&a.map[k], CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     (CUtensorMapSwizzle)swz, (CUtensorMapL2promotion)promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode %d: %d\n", k, (int)r);
    if (k == 0) { const unsigned long long *q8 = (const unsigned long long *)&a.map[0]; for (int i = 0; i < 16; i++) printf("%016llx%s", q8[i], i % 4 == 3 ? "\n" : " "); printf("d=%p\n", (void*)d); }
  }
  CUtensorMap *gm;
  cudaMalloc(&gm, sizeof(CUtensorMap));
  cudaMemcpy(gm, &a.map[0], sizeof(CUtensorMap), cudaMemcpyHostToDevice);
  a.gmap = gm;
  a.out = out;
  a.src = d;
  a.x = 37;
  a.y = 5;
  a.variant = variant;
  a.bytes = bw * bh * 2;
  a.bw = bw;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
  probe<<<1, 160, 16384>>>(a);
  e = cudaDeviceSynchronize();
  printf("variant %d sync: %s\n", variant, cudaGetErrorString(e));
  uint16_t res[240];
  cudaMemcpy(res, out, 480, cudaMemcpyDeviceToHost);
  printf("row0: %d %d %d ; row1: %d %d (want %d %d %d ; %d %d)\n", res[0], res[1], res[2], res[bw], res[bw + 1], 500 + 37, 500 + 38, 500 + 39,
         600 + 37, 600 + 38);
  return 0;
}
