// Op-level fp32 views of the projection / gather path in the reference's own tensor layouts
// (compute_pixel_in_image_mv, get_projection_features_mv, tensorflow_addons interpolate_bilinear,
// position_encoding; nerf_utils.py:64-126, 277-285).  The production path -- spatially binned
// gather / scatter through shared-memory tiles -- lives in gather_binned.cu.
#include "common.cuh"
#include "project.cuh"

namespace cn {

// ------------------------------------------------------------------------------------------
// fp32 debug view (parity tests): reference tensor layouts
// ------------------------------------------------------------------------------------------
// one texel of tf.image.resize(x, 2x, 'bilinear', half_pixel_centers) (Keras UpSampling2D(2, 'bilinear')):
//   in = (o + 0.5) / 2 - 0.5, lower = max(floor(in), 0), upper = min(ceil(in), size - 1), lerp = in - floor(in)
//   top = TL + (TR - TL) * x_lerp, bottom = BL + (BR - BL) * x_lerp, out = top + (bottom - top) * y_lerp
__device__ __forceinline__ float upsampled_texel(const float *__restrict__ x, int Hl, int Wl, int oy, int ox, int c) {
  const float iy = (oy + 0.5f) * 0.5f - 0.5f, ix = (ox + 0.5f) * 0.5f - 0.5f;   // exact in fp32
  const float fy = floorf(iy), fx = floorf(ix);
  const int y0 = max((int)fy, 0), y1 = min((int)ceilf(iy), Hl - 1);
  const int x0 = max((int)fx, 0), x1 = min((int)ceilf(ix), Wl - 1);
  const float wy = iy - fy, wx = ix - fx;
  const float tl = x[((int64_t)y0 * Wl + x0) * CN_FEAT_C + c], tr = x[((int64_t)y0 * Wl + x1) * CN_FEAT_C + c];
  const float bl = x[((int64_t)y1 * Wl + x0) * CN_FEAT_C + c], br = x[((int64_t)y1 * Wl + x1) * CN_FEAT_C + c];
  const float top = __fadd_rn(tl, __fmul_rn(__fsub_rn(tr, tl), wx));
  const float bot = __fadd_rn(bl, __fmul_rn(__fsub_rn(br, bl), wx));
  return __fadd_rn(top, __fmul_rn(__fsub_rn(bot, top), wy));
}

__global__ void __launch_bounds__(256)
project_gather_debug_kernel(CnScene sc, const float *__restrict__ ray_o, const float *__restrict__ ray_d,
                            const float *__restrict__ z, int nr, int S, float *__restrict__ pix,
                            int32_t *__restrict__ corner, float *__restrict__ gathered, float *__restrict__ cam) {
  const int64_t n_samples = (int64_t)sc.b * nr * S;
  const int64_t n = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (n >= n_samples) return;
  const int l = lane_id();
  const int64_t ray = n / S;
  const int i = (int)(n % S);
  const int b = (int)(ray / nr);
  const int r = (int)(ray % nr);
  const RayCtx rc = load_ray(ray_o, ray_d, ray, z[n]);
  const int H = sc.H, W = sc.W;
  for (int v = 0; v < sc.nv; ++v) {
    const int bv = b * sc.nv + v;
    const float *E = sc.extr_inv + (int64_t)bv * 16;
    const float *K = sc.intrinsics + (int64_t)bv * 16;
    const Proj pr = project(E, K, rc.x[0], rc.x[1], rc.x[2], H, W);
    const int64_t o = (((int64_t)bv * nr + r) * S + i);
    if (l == 0) {
      pix[o * 2 + 0] = pr.px; pix[o * 2 + 1] = pr.py;
      corner[o * 2 + 0] = pr.x0; corner[o * 2 + 1] = pr.y0;
      if (cam) {
        cam[o * 4 + 0] = pr.p[0]; cam[o * 4 + 1] = pr.p[1]; cam[o * 4 + 2] = pr.p[2];
        cam[o * 4 + 3] = dot4_rn(E + 12, rc.x[0], rc.x[1], rc.x[2], 1.f);
      }
    }
    if (!gathered) continue;
    const int64_t pix00 = ((int64_t)bv * H + pr.y0) * W + pr.x0;
    if (sc.feat_up == 2) {
      // the reference's own two-stage form, texel by texel: UpSampling2D(2, 'bilinear') of the low-resolution map at
      // the four corners, then interpolate_bilinear -- the fp32 statement the folded stencil of gather_binned.cu
      // is compared with
      const float *fl = sc.features + (int64_t)bv * (H / 2) * (W / 2) * CN_FEAT_C;
      for (int c = l; c < CN_FEAT_C; c += 32) {
        const float tl = upsampled_texel(fl, H / 2, W / 2, pr.y0, pr.x0, c), tr = upsampled_texel(fl, H / 2, W / 2, pr.y0, pr.x0 + 1, c);
        const float bl = upsampled_texel(fl, H / 2, W / 2, pr.y0 + 1, pr.x0, c), br = upsampled_texel(fl, H / 2, W / 2, pr.y0 + 1, pr.x0 + 1, c);
        gathered[o * 259 + 3 + c] = lerp2_rn(tl, tr, bl, br, pr.ax, pr.ay);
      }
    } else {
      const float *f00 = sc.features + pix00 * CN_FEAT_C;
      const float *f10 = f00 + (int64_t)W * CN_FEAT_C;
      for (int c = l; c < CN_FEAT_C; c += 32)
        gathered[o * 259 + 3 + c] = lerp2_rn(f00[c], f00[CN_FEAT_C + c], f10[c], f10[CN_FEAT_C + c], pr.ax, pr.ay);
    }
    if (l < 3) {
      const float *i00 = sc.images + pix00 * 3;
      const float *i10 = i00 + (int64_t)W * 3;
      const float tl = __fsub_rn(__fmul_rn(i00[l], 2.f), 1.f), tr = __fsub_rn(__fmul_rn(i00[3 + l], 2.f), 1.f);
      const float bl = __fsub_rn(__fmul_rn(i10[l], 2.f), 1.f), br = __fsub_rn(__fmul_rn(i10[3 + l], 2.f), 1.f);
      gathered[o * 259 + l] = lerp2_rn(tl, tr, bl, br, pr.ax, pr.ay);
    }
  }
}

// tensorflow_addons.image.interpolate_bilinear(grid [B,H,W,C], query [B,N,2], indexing='xy'):
// one warp per query, lanes stride over the channels
__global__ void __launch_bounds__(256)
interpolate_bilinear_kernel(const float *__restrict__ grid, const float *__restrict__ query, int B, int H, int W,
                            int C, int64_t N, float *__restrict__ out) {
  const int64_t n = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (n >= (int64_t)B * N) return;
  const int l = lane_id();
  const int64_t b = n / N;
  int x0, y0;
  float ax, ay, tx, ty;
  bilinear_axis(query[n * 2 + 0], W, x0, ax, tx);
  bilinear_axis(query[n * 2 + 1], H, y0, ay, ty);
  const float *g00 = grid + (((int64_t)b * H + y0) * W + x0) * C;
  const float *g10 = g00 + (int64_t)W * C;
  for (int c = l; c < C; c += 32) out[n * C + c] = lerp2_rn(g00[c], g00[C + c], g10[c], g10[C + c], ax, ay);
}

// position_encoding (nerf_utils.py:108-126): [n, 3] -> [n, 3 * n_freq * 2], 'd n f' order
__global__ void position_encoding_kernel(const float *__restrict__ pos, int64_t n, int D, int n_freq, float freq0,
                                         float *__restrict__ out) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int per = D * n_freq;
  if (idx >= n * per) return;
  const int64_t row = idx / per;
  const int c = (int)(idx % per) / n_freq, k = (int)(idx % per) % n_freq;
  const float fk = __fmul_rn(freq0, exp2f((float)k));
  float s, co;
  sincosf(__fmul_rn(pos[row * D + c], fk), &s, &co);
  out[idx * 2 + 0] = s;
  out[idx * 2 + 1] = co;
}

int check_scene(const CnScene *s, const char *who) {
  CN_REQUIRE(s, "%s: null scene", who);
  CN_REQUIRE(s->images && s->features && s->intrinsics && s->extr_inv, "%s: null scene pointer", who);
  CN_REQUIRE(s->b > 0 && s->nv > 0 && s->H >= 2 && s->W >= 2, "%s: bad scene dims b=%d nv=%d H=%d W=%d", who, s->b,
             s->nv, s->H, s->W);
  CN_REQUIRE(((uintptr_t)s->features & 15) == 0, "%s: features must be 16-byte aligned", who);
  CN_REQUIRE(s->feat_up == 1 || (s->feat_up == 2 && s->H % 2 == 0 && s->W % 2 == 0),
             "%s: feat_up must be 1, or 2 with even H and W (feat_up=%d H=%d W=%d)", who, s->feat_up, s->H, s->W);
  return 0;
}

}  // namespace cn

extern "C" int cn_project_gather_debug(const CnScene *scene, const float *ray_o, const float *ray_d, const float *z,
                                       int32_t nr, int32_t S, float *pix, int32_t *corner, float *gathered,
                                       float *cam, void *stream) {
  using namespace cn;
  if (int e = check_scene(scene, "cn_project_gather_debug")) return e;
  CN_REQUIRE(ray_o && ray_d && z && pix && corner, "cn_project_gather_debug: null pointer");
  CN_REQUIRE(nr > 0 && S > 0, "cn_project_gather_debug: nr=%d S=%d", nr, S);
  const int64_t n = (int64_t)scene->b * nr * S;
  project_gather_debug_kernel<<<(unsigned)((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(*scene, ray_o, ray_d, z, nr,
                                                                                        S, pix, corner, gathered, cam);
  return check_launch("cn_project_gather_debug");
}

extern "C" int cn_interpolate_bilinear(const float *grid, const float *query, int32_t B, int32_t H, int32_t W,
                                       int32_t C, int64_t N, float *out, void *stream) {
  using namespace cn;
  CN_REQUIRE(grid && query && out, "cn_interpolate_bilinear: null pointer");
  CN_REQUIRE(B > 0 && H >= 2 && W >= 2 && C > 0 && N >= 0, "cn_interpolate_bilinear: B=%d H=%d W=%d C=%d N=%lld", B, H,
             W, C, (long long)N);
  if (N == 0) return 0;
  const int64_t n = (int64_t)B * N;
  interpolate_bilinear_kernel<<<(unsigned)((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(grid, query, B, H, W, C, N,
                                                                                        out);
  return check_launch("cn_interpolate_bilinear");
}

extern "C" int cn_position_encoding(const float *pos, int64_t n, int32_t D, int32_t n_freq, float freq0, float *out,
                                    void *stream) {
  using namespace cn;
  CN_REQUIRE(pos && out, "cn_position_encoding: null pointer");
  CN_REQUIRE(n >= 0 && D > 0 && n_freq > 0 && n_freq <= 24, "cn_position_encoding: n=%lld D=%d n_freq=%d",
             (long long)n, D, n_freq);
  const int64_t total = n * D * n_freq;
  if (total == 0) return 0;
  position_encoding_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pos, n, D, n_freq,
                                                                                             freq0, out);
  return check_launch("cn_position_encoding");
}

