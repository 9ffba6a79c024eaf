// instantiations of the 2-D forward transform kernels: sizes with both dims <= 16
#include "tx2d.cuh"
namespace tx {
TX_SIZE_TABLE(k_tx_4x4, 4, 4)
TX_SIZE_TABLE(k_tx_8x8, 8, 8)
TX_SIZE_TABLE(k_tx_16x16, 16, 16)
TX_SIZE_TABLE(k_tx_4x8, 4, 8)
TX_SIZE_TABLE(k_tx_8x4, 8, 4)
TX_SIZE_TABLE(k_tx_8x16, 8, 16)
TX_SIZE_TABLE(k_tx_16x8, 16, 8)
TX_SIZE_TABLE(k_tx_4x16, 4, 16)
TX_SIZE_TABLE(k_tx_16x4, 16, 4)
}  // namespace tx
