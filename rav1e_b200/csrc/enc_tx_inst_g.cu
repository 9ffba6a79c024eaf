// instantiations of the fused encode_tx_block kernels with a 64-point side: 32x64, 64x32
#include "enc_tx64.cuh"
namespace enc {
ENC64_SIZE_TABLE(k_enc_32x64, 32, 64)
ENC64_SIZE_TABLE(k_enc_64x32, 64, 32)
}  // namespace enc
