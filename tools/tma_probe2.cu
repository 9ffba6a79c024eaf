// tma_probe2.cu -- Triton-shaped minimal 2-D TMA load (debugging aid)
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(const __grid_constant__ CUtensorMap map, uint16_t *out, int x, int y, int bytes, int flags) {
  __shared__ __align__(1024) unsigned char smem[8192];
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&bar)), "r"(1));
    if (flags & 1) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s32(smem)),
                 "l"((unsigned long long)&map), "r"(x), "r"(y), "r"(s32(&bar))
                 : "memory");
  }
  __syncthreads();
  asm volatile(
      "{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(s32(&bar)), "r"(0)
      : "memory");
  for (int i = threadIdx.x; i < bytes / 2; i += blockDim.x) out[i] = ((uint16_t *)smem)[i];
}
__global__ void probeD(const __grid_constant__ CUtensorMap map, uint16_t *out, int x, int y, int bytes, int flags) {
  __shared__ __align__(1024) unsigned char smem[8192];
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&bar)), "r"(32));
    if (flags & 1) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x < 32) {
   if (threadIdx.x != 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(&bar)) : "memory");
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s32(smem)),
                 "l"((unsigned long long)&map), "r"(x), "r"(y), "r"(s32(&bar))
                 : "memory");
  }
  __syncthreads();
  asm volatile(
      "{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(s32(&bar)), "r"(0)
      : "memory");
  for (int i = threadIdx.x; i < bytes / 2; i += blockDim.x) out[i] = ((uint16_t *)smem)[i];
}
struct ArgsB { alignas(64) CUtensorMap map[2]; uint16_t *out; int x, y, bytes, flags; };
__global__ void probeB(const __grid_constant__ ArgsB a) {
  __shared__ __align__(1024) unsigned char smem[8192];
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&bar)), "r"(1));
    if (a.flags & 1) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar)), "r"(a.bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s32(smem)),
                 "l"((unsigned long long)&a.map[0]), "r"(a.x), "r"(a.y), "r"(s32(&bar))
                 : "memory");
  }
  __syncthreads();
  asm volatile(
      "{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(s32(&bar)), "r"(0)
      : "memory");
  for (int i = threadIdx.x; i < a.bytes / 2; i += blockDim.x) a.out[i] = ((uint16_t *)smem)[i];
}
__global__ void probeC(const __grid_constant__ CUtensorMap map, uint16_t *out, int x, int y, int bytes, int flags) {
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t &bar = *(uint64_t *)(smem + 8192);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&bar)), "r"(1));
    if (flags & 1) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s32(smem)),
                 "l"((unsigned long long)&map), "r"(x), "r"(y), "r"(s32(&bar))
                 : "memory");
  }
  __syncthreads();
  asm volatile(
      "{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(s32(&bar)), "r"(0)
      : "memory");
  for (int i = threadIdx.x; i < bytes / 2; i += blockDim.x) out[i] = ((uint16_t *)smem)[i];
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char **argv) {
  const int flags = argc > 1 ? atoi(argv[1]) : 0;
  const int X = argc > 2 ? atoi(argv[2]) : 32;
  const int W = 4096, H = 256;
  uint16_t *h = (uint16_t *)malloc(W * H * 2), *d, *out;
  for (int i = 0; i < W * H; i++) h[i] = (uint16_t)((i / W) * 100 + (i % W) % 100);
  cudaMalloc(&d, W * H * 2);
  cudaMalloc(&out, 8192);
  cudaMemcpy(d, h, W * H * 2, cudaMemcpyHostToDevice);
  void *fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  alignas(64) CUtensorMap map;
  cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H}, strides[1] = {(cuuint64_t)W * 2};
  cuuint32_t box[2] = {16, 16}, es[2] = {1, 1};
  CUresult r = ((EncodeTiledFn)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode %d flags %d\n", (int)r, flags);
  if (flags & 2) { ArgsB b; b.map[0] = map; b.map[1] = map; b.out = out; b.x = 32; b.y = 5; b.bytes = 512; b.flags = flags; probeB<<<1, 128>>>(b); }
  else if (flags & 4) { cudaFuncSetAttribute(probeC, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384); probeC<<<1, 128, 16384>>>(map, out, 32, 5, 512, flags); }
  else if (flags & 16) probeD<<<1, (flags & 8) ? 160 : 128>>>(map, out, 32, 5, 512, flags);
  else probe<<<1, (flags & 8) ? 160 : 128>>>(map, out, X, 5, 512, flags);
  cudaError_t e = cudaDeviceSynchronize();
  printf("sync: %s\n", cudaGetErrorString(e));
  uint16_t res[256];
  cudaMemcpy(res, out, 512, cudaMemcpyDeviceToHost);
  printf("x=%d row0: %d %d %d ; row1: %d %d (want %d.. ; %d..)\n", X, res[0], res[1], res[2], res[16], res[17], 500 + X % 100, 600 + X % 100);
  return 0;
}
