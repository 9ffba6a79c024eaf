// instantiations of the 2-D inverse transform kernels: 64-point sizes
#include "itx2d.cuh"
namespace itx {
ITX_SIZE_TABLE(k_itx_64x64, 64, 64)
ITX_SIZE_TABLE(k_itx_32x64, 32, 64)
ITX_SIZE_TABLE(k_itx_16x64, 16, 64)
ITX_SIZE_TABLE(k_itx_64x32, 64, 32)
ITX_SIZE_TABLE(k_itx_64x16, 64, 16)
}  // namespace itx
