// Small kernels of the pose-query path (DeltaNGF / LanguageNeRF / GraspMVNeRF `_call`,
// delta_ngf/model.py:188-245): the gradient of the four tapped stage outputs of the fine embedding
// (`fine_embedding(...)[4:]`, model.py:240-241) enters the bf16 gradient stream of the block backward.
#include "common.cuh"
#include "project.cuh"

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

// Tangent of the MLP input rows of the pose-query call for a perturbation (u_p, u_d) of the query points and
// approach directions (the double backward of DeltaNGF.train_step, delta_ngf/model.py:264-296): per sample-view row
//   p' = E3 u_p, d' = E3 u_d;  q' = K3 p';  pix' = (q'.xy - pix q'.z) / max(q.z, 1e-8)  (zero where a clamp is active)
//   feature / image columns: pix'_x d(bilinear)/dx + pix'_y d(bilinear)/dy from the four corner texels
//   posenc columns: f_k p'_c (cos, -sin)(f_k p_c), same for the direction E [d; 1] (quirk Q1)
// written as bf16 MLP-input tiles in the kernels' column order [feat 256 | posenc(p) 60 | posenc(d) 60 | img 3 | 0].
// One warp per row; lane l owns feature channels 8l .. 8l+7 and (l < 16) columns 256 + 8l .. of the rest.
__global__ void __launch_bounds__(256)
query_tangent_x_kernel(CnScene sc, const float *__restrict__ pts, const float *__restrict__ dirs,
                       const float *__restrict__ u_pts, const float *__restrict__ u_dirs, int n, int64_t rows,
                       uint8_t *__restrict__ x_tiles) {
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const uint32_t l = lane_id();
  const int bv = (int)(row / n), i = (int)(row % n), b = bv / sc.nv;
  const int64_t pt = (int64_t)b * n + i;
  const float *E = sc.extr_inv + (int64_t)bv * 16, *K = sc.intrinsics + (int64_t)bv * 16;
  const float x = pts[pt * 3], y = pts[pt * 3 + 1], z = pts[pt * 3 + 2];
  const float dx = dirs[pt * 3], dy = dirs[pt * 3 + 1], dz = dirs[pt * 3 + 2];
  const float ux = u_pts[pt * 3], uy = u_pts[pt * 3 + 1], uz = u_pts[pt * 3 + 2];
  const float vx = u_dirs[pt * 3], vy = u_dirs[pt * 3 + 1], vz = u_dirs[pt * 3 + 2];
  const Proj pr = project(E, K, x, y, z, sc.H, sc.W);
  float pd[3], dd[3], dc[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    pd[c] = E[4 * c] * ux + E[4 * c + 1] * uy + E[4 * c + 2] * uz;
    dd[c] = E[4 * c] * vx + E[4 * c + 1] * vy + E[4 * c + 2] * vz;
    dc[c] = dot4_rn(E + 4 * c, dx, dy, dz, 1.f);
  }
  const float qdx = K[0] * pd[0] + K[1] * pd[1] + K[2] * pd[2];
  const float qdy = K[4] * pd[0] + K[5] * pd[1] + K[6] * pd[2];
  const float qdz = pr.q[2] >= 1e-8f ? K[8] * pd[0] + K[9] * pd[1] + K[10] * pd[2] : 0.f;
  const bool ok_x = pr.tx >= 0.f && pr.tx <= 1.f && pr.rx >= -1e6f && pr.rx <= 1e6f;
  const bool ok_y = pr.ty >= 0.f && pr.ty <= 1.f && pr.ry >= -1e6f && pr.ry <= 1e6f;
  const float pxd = ok_x ? (qdx - pr.rx * qdz) / pr.den : 0.f;
  const float pyd = ok_y ? (qdy - pr.ry * qdz) / pr.den : 0.f;
  const float ax = pr.ax, ay = pr.ay;
  const int64_t pix00 = ((int64_t)bv * sc.H + pr.y0) * sc.W + pr.x0;
  uint8_t *dst = x_tiles + (row >> 7) * (int64_t)CN_X_TILE_BYTES + (uint32_t)(row & 127) * 128u;
  const uint32_t sw = (uint32_t)row & 7u;
  auto lerp_tangent = [&](float tl, float tr, float bl, float br) {
    const float gx = (1.f - ay) * (tr - tl) + ay * (br - bl), gy = (1.f - ax) * (bl - tl) + ax * (br - tr);
    return pxd * gx + pyd * gy;
  };
  {
    const float *f00 = sc.features + pix00 * CN_FEAT_C + 8 * l, *f10 = f00 + (int64_t)sc.W * CN_FEAT_C;
    float v[8];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float4 tl = ldg_nc_f4(f00 + 4 * h), tr = ldg_nc_f4(f00 + CN_FEAT_C + 4 * h);
      const float4 bl = ldg_nc_f4(f10 + 4 * h), br = ldg_nc_f4(f10 + CN_FEAT_C + 4 * h);
      v[4 * h + 0] = lerp_tangent(tl.x, tr.x, bl.x, br.x); v[4 * h + 1] = lerp_tangent(tl.y, tr.y, bl.y, br.y);
      v[4 * h + 2] = lerp_tangent(tl.z, tr.z, bl.z, br.z); v[4 * h + 3] = lerp_tangent(tl.w, tr.w, bl.w, br.w);
    }
    *reinterpret_cast<uint4 *>(dst + (l >> 3) * CN_SLAB_BYTES + (((l & 7u) ^ sw) << 4)) =
        make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  }
  if (l < 16) {
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int j = (int)l * 8 + e;        // column 256 + j
      float val = 0.f;
      if (j < 120) {
        const int jj = j < 60 ? j : j - 60;
        const int c = jj / 20, k = (jj % 20) >> 1;
        const float f = pe_freq(k);
        const float base = j < 60 ? pr.p[c] : dc[c], tang = j < 60 ? pd[c] : dd[c];
        float sn, cs;
        sincosf(__fmul_rn(base, f), &sn, &cs);
        val = (jj & 1) ? -sn * f * tang : cs * f * tang;
      } else if (j < 123) {
        const int c = j - 120;
        const float *i00 = sc.images + pix00 * 3 + c, *i10 = i00 + (int64_t)sc.W * 3;
        val = 2.f * lerp_tangent(i00[0], i00[3], i10[0], i10[3]);
      }
      v[e] = val;
    }
    const uint32_t chunk = 32u + l;        // chunks 32..47 of the 48-chunk X row
    *reinterpret_cast<uint4 *>(dst + (chunk >> 3) * CN_SLAB_BYTES + (((chunk & 7u) ^ sw) << 4)) =
        make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  }
}

}  // namespace cn

extern "C" int cn_query_tangent_x(const CnScene *scene, const float *points, const float *dirs, const float *u_points,
                                  const float *u_dirs, int32_t n, void *x_tiles, void *stream) {
  using namespace cn;
  if (int rc = check_scene(scene, "cn_query_tangent_x")) return rc;
  CN_REQUIRE(points && dirs && u_points && u_dirs && x_tiles, "cn_query_tangent_x: null pointer");
  CN_REQUIRE(scene->feat_up == 1, "cn_query_tangent_x: full-resolution feature maps only");
  CN_REQUIRE(n > 0 && n % CN_TILE_ROWS == 0, "cn_query_tangent_x: n=%d must be a positive multiple of %d", n, CN_TILE_ROWS);
  CN_REQUIRE(((uintptr_t)x_tiles & 127) == 0, "cn_query_tangent_x: x_tiles must be 128-byte aligned");
  const int64_t rows = (int64_t)scene->b * scene->nv * n;
  query_tangent_x_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, (cudaStream_t)stream>>>(*scene, points, dirs, u_points, u_dirs,
                                                                                       n, rows, (uint8_t *)x_tiles);
  return check_launch("cn_query_tangent_x");
}

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
