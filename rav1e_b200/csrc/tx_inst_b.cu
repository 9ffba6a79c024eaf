// instantiations of the 2-D forward transform kernels: sizes with a 32-point side
#include "tx2d.cuh"
namespace tx {
TX_SIZE_TABLE(k_tx_32x32, 32, 32)
TX_SIZE_TABLE(k_tx_16x32, 16, 32)
TX_SIZE_TABLE(k_tx_32x16, 32, 16)
TX_SIZE_TABLE(k_tx_8x32, 8, 32)
TX_SIZE_TABLE(k_tx_32x8, 32, 8)
}  // namespace tx
