// instantiations of the 2-D inverse transform kernels: 32-point sizes
#include "itx2d.cuh"
namespace itx {
ITX_SIZE_TABLE(k_itx_32x32, 32, 32)
ITX_SIZE_TABLE(k_itx_16x32, 16, 32)
ITX_SIZE_TABLE(k_itx_32x16, 32, 16)
ITX_SIZE_TABLE(k_itx_8x32, 8, 32)
ITX_SIZE_TABLE(k_itx_32x8, 32, 8)
}  // namespace itx
