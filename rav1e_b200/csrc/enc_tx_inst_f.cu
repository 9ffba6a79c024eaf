// instantiations of the fused encode_tx_block kernels with a 64-point side: 64x64, 64x16, 16x64
#include "enc_tx64.cuh"
namespace enc {
ENC64_SIZE_TABLE(k_enc_64x64, 64, 64)
ENC64_SIZE_TABLE(k_enc_64x16, 64, 16)
ENC64_SIZE_TABLE(k_enc_16x64, 16, 64)
}  // namespace enc
