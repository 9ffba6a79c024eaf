// instantiations of the fused encode_tx_block kernels: 16x16, 8x16, 16x8
#include "enc_tx.cuh"
namespace enc {
ENC_SIZE_TABLE(k_enc_16x16, 16, 16)
ENC_SIZE_TABLE(k_enc_8x16, 8, 16)
ENC_SIZE_TABLE(k_enc_16x8, 16, 8)
}  // namespace enc
