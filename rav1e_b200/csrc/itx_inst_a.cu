// instantiations of the 2-D inverse transform kernels: sizes with both dims <= 16
#include "itx2d.cuh"
namespace itx {
ITX_SIZE_TABLE(k_itx_4x4, 4, 4)
ITX_SIZE_TABLE(k_itx_8x8, 8, 8)
ITX_SIZE_TABLE(k_itx_16x16, 16, 16)
ITX_SIZE_TABLE(k_itx_4x8, 4, 8)
ITX_SIZE_TABLE(k_itx_8x4, 8, 4)
ITX_SIZE_TABLE(k_itx_8x16, 8, 16)
ITX_SIZE_TABLE(k_itx_16x8, 16, 8)
ITX_SIZE_TABLE(k_itx_4x16, 4, 16)
ITX_SIZE_TABLE(k_itx_16x4, 16, 4)
}  // namespace itx
