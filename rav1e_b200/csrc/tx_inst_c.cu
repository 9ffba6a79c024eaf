// instantiations of the 2-D forward transform kernels: 64-point columns
#include "tx2d.cuh"
namespace tx {
TX_SIZE_TABLE(k_tx_64x64, 64, 64)
TX_SIZE_TABLE(k_tx_32x64, 32, 64)
TX_SIZE_TABLE(k_tx_16x64, 16, 64)
}  // namespace tx
