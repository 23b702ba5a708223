// cn_pack_weights: fp32 Keras-layout parameters of one MVResNetMLPNeRFEmbedding(6,128) +
// RenderReadout(4) (layers.py:334-397)  ->  bf16 tcgen05 operand images + fp32 bias tables.
//
// Packed layout (bytes):
//   [0, 491520)        forward B operands, K-major [n = out][k = in]:
//                        W0 as 6 slabs (input columns in the internal X order, rows 379..383 zero),
//                        then block k (0..5) layer j (0,1) at 98304 + (2k+j)*32768
//   [491520, 983040)   dgrad B operands, K-major [n = in][k = out] (i.e. the Keras matrix itself):
//                        W0 as 3 chunks of 128 input columns, then the 12 block matrices
//   [983040, 999424)   fp32 tables (float index):
//                        [0, 896)     per-view stage: F[0]=b0, F[1+2k]=ba_k, F[2+2k]=b0+sum_{i<=k} bb_i
//                        [896, 1792)  fusion stage:   U[0]=0,  U[1+2k]=ba'_k, U[2+2k]=sum_{i<=k} bb'_i
//                        [1792, 2304) readout kernel [128][4], [2304, 2308) readout bias
#include "common.cuh"
#include "mlp_layout.cuh"

namespace cn {

__global__ void pack_weights_kernel(const float *__restrict__ p, uint8_t *__restrict__ out) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nthreads = gridDim.x * blockDim.x;
  // ---- first Dense: 128 x 384 elements, both directions
  for (int e = tid; e < 128 * CN_IN_PAD; e += nthreads) {
    const int n = e / CN_IN_PAD, kint = e % CN_IN_PAD;  // n = output unit, kint = internal input column
    const int ref = x_col_to_ref(kint);
    const float w = ref >= 0 ? p[kParamW0 + ref * 128 + n] : 0.f;
    const __nv_bfloat16 wb = __float2bfloat16_rn(w);
    *reinterpret_cast<__nv_bfloat16 *>(out + kPackFwdW0 + tile_off(n, kint)) = wb;
    const int c = kint >> 7, nl = kint & 127;
    *reinterpret_cast<__nv_bfloat16 *>(out + kPackBwdW0 + c * CN_ACT_TILE_BYTES + tile_off(nl, n)) = wb;
  }
  // ---- 12 block matrices
  for (int e = tid; e < 12 * 128 * 128; e += nthreads) {
    const int m = e / 16384, in = (e / 128) % 128, o = e % 128;
    const int blk = m >> 1, j = m & 1;
    const float w = p[param_block_w(blk, j) + in * 128 + o];
    const __nv_bfloat16 wb = __float2bfloat16_rn(w);
    *reinterpret_cast<__nv_bfloat16 *>(out + pack_fwd_block(blk, j) + tile_off(o, in)) = wb;
    *reinterpret_cast<__nv_bfloat16 *>(out + pack_bwd_block(blk, j) + tile_off(in, o)) = wb;
  }
  // ---- bias tables
  float *tab = reinterpret_cast<float *>(out + kPackTables);
  for (int c = tid; c < 128; c += nthreads) {
    float cum = p[kParamB0 + c];
    tab[kTabFeature + c] = cum;
    for (int k = 0; k < 3; ++k) {
      tab[kTabFeature + (1 + 2 * k) * 128 + c] = p[param_block_b(k, 0) + c];
      cum += p[param_block_b(k, 1) + c];
      tab[kTabFeature + (2 + 2 * k) * 128 + c] = cum;
    }
    cum = 0.f;
    tab[kTabFusion + c] = 0.f;
    for (int k = 0; k < 3; ++k) {
      tab[kTabFusion + (1 + 2 * k) * 128 + c] = p[param_block_b(3 + k, 0) + c];
      cum += p[param_block_b(3 + k, 1) + c];
      tab[kTabFusion + (2 + 2 * k) * 128 + c] = cum;
    }
    for (int j = 0; j < 4; ++j) tab[kTabReadoutW + c * 4 + j] = p[kParamWr + c * 4 + j];
    if (c < 4) tab[kTabReadoutB + c] = p[kParamBr + c];
  }
}

}  // namespace cn

extern "C" int cn_pack_weights(const float *params, void *packed, void *stream) {
  using namespace cn;
  CN_REQUIRE(params && packed, "cn_pack_weights: null pointer");
  CN_REQUIRE(((uintptr_t)packed & 127) == 0, "cn_pack_weights: packed buffer must be 128-byte aligned");
  pack_weights_kernel<<<296, 256, 0, (cudaStream_t)stream>>>(params, (uint8_t *)packed);
  return check_launch("cn_pack_weights");
}
