// instantiations of the fused encode_tx_block kernels: 8x32, 32x8
#include "enc_tx.cuh"
namespace enc {
ENC_SIZE_TABLE(k_enc_8x32, 8, 32)
ENC_SIZE_TABLE(k_enc_32x8, 32, 8)
}  // namespace enc
