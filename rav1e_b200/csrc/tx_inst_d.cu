// instantiations of the 2-D forward transform kernels: 64-point rows
#include "tx2d.cuh"
namespace tx {
TX_SIZE_TABLE(k_tx_64x32, 64, 32)
TX_SIZE_TABLE(k_tx_64x16, 64, 16)
}  // namespace tx
