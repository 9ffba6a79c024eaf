// instantiations of the fused encode_tx_block kernels: 16x32, 32x16
#include "enc_tx.cuh"
namespace enc {
ENC_SIZE_TABLE(k_enc_16x32, 16, 32)
ENC_SIZE_TABLE(k_enc_32x16, 32, 16)
}  // namespace enc
