// instantiations of the fused encode_tx_block kernels: 32x32
#include "enc_tx.cuh"
namespace enc {
ENC_SIZE_TABLE(k_enc_32x32, 32, 32)
}  // namespace enc
