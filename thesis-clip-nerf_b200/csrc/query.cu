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

// poses = transforms . transforms_to_check (delta_ngf/model.py:199-200, 221-223): query point = M C_j [0 0 0 1]^T,
// approach direction = (M C_j)[:3, :3] z, rows ordered '(np n5)'.  The bottom row of M is [0 0 0 1], so
// point = R c_j + t and direction = R (C_j[:3, 2]).  One thread per (pose, offset).
__global__ void __launch_bounds__(256) pose_points_fwd_kernel(const float *__restrict__ M, const float *__restrict__ ttc,
                                                              int64_t n_poses, int n5, float *__restrict__ pts,
                                                              float *__restrict__ dirs) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n_poses * n5) return;
  const int64_t pose = i / n5;
  const int j = (int)(i % n5);
  const float *m = M + pose * 16, *c = ttc + j * 16;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    // same association as the broadcast product + sum of the torch form: ((m0 c0 + m1 c1) + m2 c2) + t
    const float p = (m[r * 4 + 0] * c[3] + m[r * 4 + 1] * c[7]) + m[r * 4 + 2] * c[11];
    pts[i * 3 + r] = p + m[r * 4 + 3];
    dirs[i * 3 + r] = (m[r * 4 + 0] * c[2] + m[r * 4 + 1] * c[6]) + m[r * 4 + 2] * c[10];
  }
}
// d M[pose] = sum_j d_pt c_j^T + d_dir z_j^T (rotation part), sum_j d_pt (translation); one thread per pose
__global__ void __launch_bounds__(128) pose_points_bwd_kernel(const float *__restrict__ d_pts, const float *__restrict__ d_dirs,
                                                              const float *__restrict__ ttc, int64_t n_poses, int n5,
                                                              float *__restrict__ dM) {
  const int64_t pose = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (pose >= n_poses) return;
  float g[12] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int j = 0; j < n5; ++j) {
    const float *c = ttc + j * 16;
    const float *dp = d_pts + (pose * n5 + j) * 3, *dd = d_dirs + (pose * n5 + j) * 3;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const float a = dp[r], b = dd[r];
      g[r * 4 + 0] += a * c[3] + b * c[2];
      g[r * 4 + 1] += a * c[7] + b * c[6];
      g[r * 4 + 2] += a * c[11] + b * c[10];
      g[r * 4 + 3] += a;
    }
  }
#pragma unroll
  for (int k = 0; k < 12; ++k) dM[pose * 16 + k] = g[k];
#pragma unroll
  for (int k = 12; k < 16; ++k) dM[pose * 16 + k] = 0.f;
}

}  // namespace cn

extern "C" int cn_pose_points_fwd(const float *transforms, const float *transforms_to_check, int64_t n_poses, int32_t n5,
                                  float *points, float *dirs, void *stream) {
  using namespace cn;
  CN_REQUIRE(transforms && transforms_to_check && points && dirs, "cn_pose_points_fwd: null pointer");
  CN_REQUIRE(n_poses >= 0 && n5 > 0, "cn_pose_points_fwd: n_poses=%lld n5=%d", (long long)n_poses, n5);
  if (n_poses == 0) return 0;
  const int64_t n = n_poses * n5;
  pose_points_fwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(transforms, transforms_to_check,
                                                                                        n_poses, n5, points, dirs);
  return check_launch("cn_pose_points_fwd");
}

extern "C" int cn_pose_points_bwd(const float *d_points, const float *d_dirs, const float *transforms_to_check,
                                  int64_t n_poses, int32_t n5, float *d_transforms, void *stream) {
  using namespace cn;
  CN_REQUIRE(d_points && d_dirs && transforms_to_check && d_transforms, "cn_pose_points_bwd: null pointer");
  CN_REQUIRE(n_poses >= 0 && n5 > 0, "cn_pose_points_bwd: n_poses=%lld n5=%d", (long long)n_poses, n5);
  if (n_poses == 0) return 0;
  pose_points_bwd_kernel<<<(unsigned)((n_poses + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      d_points, d_dirs, transforms_to_check, n_poses, n5, d_transforms);
  return check_launch("cn_pose_points_bwd");
}

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
