// Offsets of the flat fp32 parameter vector and of the packed operand buffer of one net
// (see cn_pack_weights in include/clipnerf_b200.h and pack.cu).
#pragma once
#include "common.cuh"

namespace cn {

// ---- flat fp32 parameters (float offsets), Keras variable order, kernel[in][out]
constexpr int kParamW0 = 0;                       // layer_0.kernel [379][128]
constexpr int kParamB0 = CN_IN_DIM * 128;         // layer_0.bias
constexpr int kParamBlocks = kParamB0 + 128;      // 6 blocks x (Wa, ba, Wb, bb)
constexpr int kParamBlockStride = 2 * (128 * 128 + 128);
constexpr int kParamWr = CN_EMB_PARAMS;           // output_layer.kernel [128][4]
constexpr int kParamBr = CN_EMB_PARAMS + 512;     // output_layer.bias [4]
static_assert(kParamBlocks + 6 * kParamBlockStride == CN_EMB_PARAMS, "parameter count");
__host__ __device__ constexpr int param_block_w(int blk, int j) {
  return kParamBlocks + blk * kParamBlockStride + j * (128 * 128 + 128);
}
__host__ __device__ constexpr int param_block_b(int blk, int j) { return param_block_w(blk, j) + 128 * 128; }

// ---- packed buffer (byte offsets)
constexpr int kPackFwdW0 = 0;
constexpr int kPackFwdBlocks = CN_X_TILE_BYTES;
constexpr int kPackBwd = 491520;
constexpr int kPackBwdW0 = kPackBwd;
constexpr int kPackBwdBlocks = kPackBwd + 3 * CN_ACT_TILE_BYTES;
constexpr int kPackTables = 2 * 491520;
static_assert(kPackFwdBlocks + 12 * CN_ACT_TILE_BYTES == kPackBwd, "packed layout");
static_assert(kPackTables + 16384 == CN_PACKED_BYTES, "packed layout");
__host__ __device__ constexpr int pack_fwd_block(int blk, int j) {
  return kPackFwdBlocks + (2 * blk + j) * CN_ACT_TILE_BYTES;
}
__host__ __device__ constexpr int pack_bwd_block(int blk, int j) {
  return kPackBwdBlocks + (2 * blk + j) * CN_ACT_TILE_BYTES;
}
// fp32 tables (float offsets inside the table region)
constexpr int kTabFeature = 0;       // 7 x 128
constexpr int kTabFusion = 896;      // 7 x 128
constexpr int kTabReadoutW = 1792;   // [128][4]
constexpr int kTabReadoutB = 2304;   // [4]
constexpr int kTabFloats = 2308;

}  // namespace cn
