// tma_probe3.cu -- which shared-memory alignment does a 2-D tensor load need on this part?  (debugging aid)
// usage: tma_probe3 <dst offset in bytes>   -- loads a 24 x 11 u16 box to smem + offset and checks the contents
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(const __grid_constant__ CUtensorMap map, uint16_t *out, int x, int y, int bytes, int off) {
  __shared__ __align__(1024) unsigned char smem[8192];
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&bar)), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s32(smem + off)),
                 "l"((unsigned long long)&map), "r"(x), "r"(y), "r"(s32(&bar))
                 : "memory");
  }
  __syncthreads();
  asm volatile(
      "{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(s32(&bar)), "r"(0)
      : "memory");
  for (int i = threadIdx.x; i < bytes / 2; i += blockDim.x) out[i] = ((uint16_t *)(smem + off))[i];
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char **argv) {
  const int off = argc > 1 ? atoi(argv[1]) : 0;
  const int W = 4096, H = 256, BW = 24, BH = 11;
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
  cuuint32_t box[2] = {BW, BH}, es[2] = {1, 1};
  CUresult r = ((EncodeTiledFn)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  probe<<<1, 128>>>(map, out, 32, 5, BW * BH * 2, off);
  cudaError_t e = cudaDeviceSynchronize();
  uint16_t res[BW * BH];
  int bad = 0;
  if (e == cudaSuccess) {
    cudaMemcpy(res, out, sizeof res, cudaMemcpyDeviceToHost);
    for (int y = 0; y < BH; y++)
      for (int x = 0; x < BW; x++) bad += res[y * BW + x] != h[(5 + y) * W + 32 + x];
  }
  printf("dst offset %d: encode %d, sync: %s, mismatches %d\n", off, (int)r, cudaGetErrorString(e), bad);
  return 0;
}
