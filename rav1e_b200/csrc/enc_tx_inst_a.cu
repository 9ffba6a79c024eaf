// instantiations of the fused encode_tx_block kernels: 4x4, 8x8, 4x8, 8x4, 4x16, 16x4
#include "enc_tx.cuh"
namespace enc {
ENC_SIZE_TABLE(k_enc_4x4, 4, 4)
ENC_SIZE_TABLE(k_enc_8x8, 8, 8)
ENC_SIZE_TABLE(k_enc_4x8, 4, 8)
ENC_SIZE_TABLE(k_enc_8x4, 8, 4)
ENC_SIZE_TABLE(k_enc_4x16, 4, 16)
ENC_SIZE_TABLE(k_enc_16x4, 16, 4)
}  // namespace enc
