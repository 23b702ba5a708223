// Small kernels of the pose-query path (DeltaNGF / LanguageNeRF / GraspMVNeRF `_call`,
// delta_ngf/model.py:188-245): the gradient of the four tapped stage outputs of the fine embedding
// (`fine_embedding(...)[4:]`, model.py:240-241) enters the bf16 gradient stream of the block backward.
#include "common.cuh"

namespace cn {

// g_tiles[tile][row][col] (bf16, swizzled activation-tile image) (+)= rows[tile*128 + row][col] (fp32, row-major).
// One thread per 16-byte chunk (8 columns): two coalesced 16-byte fp32 loads, one 16-byte store.
__global__ void __launch_bounds__(256) tile_add_rows_kernel(uint8_t *__restrict__ g_tiles,
                                                            const float *__restrict__ rows, int64_t n_chunks,
                                                            int accumulate) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n_chunks) return;
  const int64_t grow = i >> 4;             // global row
  const uint32_t chunk = (uint32_t)(i & 15);
  const int64_t tile = grow >> 7;
  const uint32_t row = (uint32_t)(grow & 127);
  const float4 a = ldg_nc_f4(rows + i * 8);
  const float4 b = ldg_nc_f4(rows + i * 8 + 4);
  float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  uint4 *dst = reinterpret_cast<uint4 *>(g_tiles + tile * (int64_t)CN_ACT_TILE_BYTES + tile_chunk_off(row, chunk));
  if (accumulate) {
    const uint4 o = *dst;
    v[0] += bf16_lo(o.x); v[1] += bf16_hi(o.x); v[2] += bf16_lo(o.y); v[3] += bf16_hi(o.y);
    v[4] += bf16_lo(o.z); v[5] += bf16_hi(o.z); v[6] += bf16_lo(o.w); v[7] += bf16_hi(o.w);
  }
  *dst = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
}

}  // namespace cn

extern "C" int cn_tile_add_rows(void *g_tiles, const float *rows, int64_t n_tiles, int32_t accumulate, void *stream) {
  using namespace cn;
  CN_REQUIRE(g_tiles && rows, "cn_tile_add_rows: null pointer");
  CN_REQUIRE(n_tiles >= 0, "cn_tile_add_rows: n_tiles %lld", (long long)n_tiles);
  CN_REQUIRE((((uintptr_t)g_tiles | (uintptr_t)rows) & 15) == 0, "cn_tile_add_rows: buffers must be 16-byte aligned");
  if (n_tiles == 0) return 0;
  const int64_t n_chunks = n_tiles * 128 * 16;
  tile_add_rows_kernel<<<(unsigned)((n_chunks + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (uint8_t *)g_tiles, rows, n_chunks, accumulate);
  return check_launch("cn_tile_add_rows");
}
