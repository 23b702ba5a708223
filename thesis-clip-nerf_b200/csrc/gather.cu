// Projection of ray samples into the source views, edge-clamped bilinear gather of the
// 256-channel feature map + 3-channel image, positional encoding, and assembly of the MLP
// input rows -- the fused replacement of compute_pixel_in_image_mv, get_projection_features_mv,
// world_to_camera_direction_vector_mv, position_encoding and the concat of layers.py:362
// (nerf_utils.py:64-126, 277-285).  The 259-channel concat copy of the reference
// (nerf_utils.py:278) does not exist here: image and feature map are read in place.
//
// One warp per ray sample; the warp loops over the source views.  Per view and corner the warp
// reads one 1 KiB feature row with two fully coalesced 512-byte requests (float4 per lane), and
// writes the bf16 MLP-input row as 128-byte segments of the swizzled tile image.
#include "common.cuh"

namespace cn {

struct Proj {
  float p[3];    // point in the source camera frame (MLP position input, quirk Q2)
  float q[3];    // K . p
  float den;     // max(q.z, 1e-8)
  float rx, ry;  // unclipped ratios
  float px, py;  // clipped pixel coordinates
  int x0, y0;    // clamped floor (top-left corner)
  float ax, ay;  // clamped interpolation weights
  float tx, ty;  // unclamped fractional offsets (pix - floor)
};

// Explicitly rounded, left-to-right products/sums: the oracle pins the same order, which makes
// pixel coordinates and corner indices bit-exact.
__device__ __forceinline__ float dot4_rn(const float *m, float x, float y, float z, float w) {
  float acc = __fmul_rn(m[0], x);
  acc = __fadd_rn(acc, __fmul_rn(m[1], y));
  acc = __fadd_rn(acc, __fmul_rn(m[2], z));
  acc = __fadd_rn(acc, __fmul_rn(m[3], w));
  return acc;
}

__device__ __forceinline__ void bilinear_axis(float q, int size, int &f0, float &alpha, float &t) {
  float fl = floorf(q);
  fl = fminf(fmaxf(fl, 0.f), (float)(size - 2));
  f0 = (int)fl;
  t = __fsub_rn(q, fl);
  alpha = fminf(fmaxf(t, 0.f), 1.f);
}

__device__ __forceinline__ Proj project(const float *__restrict__ E, const float *__restrict__ K, float x, float y,
                                        float z, int H, int W) {
  Proj r;
  float p3;
  r.p[0] = dot4_rn(E + 0, x, y, z, 1.f);
  r.p[1] = dot4_rn(E + 4, x, y, z, 1.f);
  r.p[2] = dot4_rn(E + 8, x, y, z, 1.f);
  p3 = dot4_rn(E + 12, x, y, z, 1.f);
  r.q[0] = dot4_rn(K + 0, r.p[0], r.p[1], r.p[2], p3);
  r.q[1] = dot4_rn(K + 4, r.p[0], r.p[1], r.p[2], p3);
  r.q[2] = dot4_rn(K + 8, r.p[0], r.p[1], r.p[2], p3);
  r.den = fmaxf(r.q[2], 1e-8f);
  r.rx = __fdiv_rn(r.q[0], r.den);
  r.ry = __fdiv_rn(r.q[1], r.den);
  r.px = fminf(fmaxf(r.rx, -1e6f), 1e6f);
  r.py = fminf(fmaxf(r.ry, -1e6f), 1e6f);
  bilinear_axis(r.px, W, r.x0, r.ax, r.tx);
  bilinear_axis(r.py, H, r.y0, r.ay, r.ty);
  return r;
}

__device__ __forceinline__ float lerp2_rn(float tl, float tr, float bl, float br, float ax, float ay) {
  const float top = __fadd_rn(__fmul_rn(ax, __fsub_rn(tr, tl)), tl);
  const float bot = __fadd_rn(__fmul_rn(ax, __fsub_rn(br, bl)), bl);
  return __fadd_rn(__fmul_rn(ay, __fsub_rn(bot, top)), top);
}
__device__ __forceinline__ float4 lerp2_f4(float4 tl, float4 tr, float4 bl, float4 br, float ax, float ay) {
  return make_float4(lerp2_rn(tl.x, tr.x, bl.x, br.x, ax, ay), lerp2_rn(tl.y, tr.y, bl.y, br.y, ax, ay),
                     lerp2_rn(tl.z, tr.z, bl.z, br.z, ax, ay), lerp2_rn(tl.w, tr.w, bl.w, br.w, ax, ay));
}

__device__ __forceinline__ float pe_freq(int k) { return __fmul_rn(3.14159274101257324f, (float)(1 << k)); }

struct RayCtx {
  float o[3], d[3], x[3];
};
__device__ __forceinline__ RayCtx load_ray(const float *__restrict__ ray_o, const float *__restrict__ ray_d,
                                           int64_t ray, float zz) {
  RayCtx c;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    c.o[i] = ray_o[ray * 3 + i];
    c.d[i] = ray_d[ray * 3 + i];
    c.x[i] = __fadd_rn(c.o[i], __fmul_rn(zz, c.d[i]));  // world point o + z d
  }
  return c;
}

// ------------------------------------------------------------------------------------------
// forward: bf16 tile images
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
project_gather_fwd_kernel(CnScene sc, const float *__restrict__ ray_o, const float *__restrict__ ray_d,
                          const float *__restrict__ z, int nr, int S, uint8_t *__restrict__ x_tiles) {
  const int64_t n_samples = (int64_t)sc.b * nr * S;
  const int64_t n = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (n >= n_samples) return;
  const int l = lane_id();
  const int64_t ray = n / S;  // (b, r)
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
    const int64_t pix00 = ((int64_t)bv * H + pr.y0) * W + pr.x0;
    const float *f00 = sc.features + pix00 * CN_FEAT_C;
    const float *f10 = f00 + (int64_t)W * CN_FEAT_C;
    // 8 independent 16-byte loads per lane: 4 corners x 2 halves of the 256-channel row
    const float4 tl0 = ldg_nc_f4(f00 + 4 * l), tl1 = ldg_nc_f4(f00 + 128 + 4 * l);
    const float4 tr0 = ldg_nc_f4(f00 + CN_FEAT_C + 4 * l), tr1 = ldg_nc_f4(f00 + CN_FEAT_C + 128 + 4 * l);
    const float4 bl0 = ldg_nc_f4(f10 + 4 * l), bl1 = ldg_nc_f4(f10 + 128 + 4 * l);
    const float4 br0 = ldg_nc_f4(f10 + CN_FEAT_C + 4 * l), br1 = ldg_nc_f4(f10 + CN_FEAT_C + 128 + 4 * l);

    const int64_t row = ((int64_t)bv * nr + r) * S + i;
    uint8_t *tile = x_tiles + (row >> 7) * (int64_t)CN_X_TILE_BYTES;
    const uint32_t rr = (uint32_t)(row & 127);

    // positional encodings (lanes 0..29: one (coordinate, frequency) pair each), image (lane 31)
    if (l < 30) {
      const int c = l / 10, k = l % 10;
      const float fk = pe_freq(k);
      float s, co;
      sincosf(__fmul_rn(pr.p[c], fk), &s, &co);
      *reinterpret_cast<uint32_t *>(tile + tile_off(rr, 256 + 2 * l)) = pack_bf16x2(s, co);
      // direction in the source camera frame with homogeneous w = 1 (quirk Q1)
      const float dc = dot4_rn(E + 4 * c, rc.d[0], rc.d[1], rc.d[2], 1.f);
      sincosf(__fmul_rn(dc, fk), &s, &co);
      *reinterpret_cast<uint32_t *>(tile + tile_off(rr, 316 + 2 * l)) = pack_bf16x2(s, co);
    } else if (l == 31) {
      const float *i00 = sc.images + pix00 * 3;
      const float *i10 = i00 + (int64_t)W * 3;
      float g[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float tl = __fsub_rn(__fmul_rn(i00[c], 2.f), 1.f), tr = __fsub_rn(__fmul_rn(i00[3 + c], 2.f), 1.f);
        const float bl = __fsub_rn(__fmul_rn(i10[c], 2.f), 1.f), br = __fsub_rn(__fmul_rn(i10[3 + c], 2.f), 1.f);
        g[c] = lerp2_rn(tl, tr, bl, br, pr.ax, pr.ay);
      }
      uint4 o = make_uint4(pack_bf16x2(g[0], g[1]), pack_bf16x2(g[2], 0.f), 0u, 0u);
      *reinterpret_cast<uint4 *>(tile + tile_chunk_off(rr, 47)) = o;
    }
    const float4 g0 = lerp2_f4(tl0, tr0, bl0, br0, pr.ax, pr.ay);
    const float4 g1 = lerp2_f4(tl1, tr1, bl1, br1, pr.ax, pr.ay);
    // channels 4l..4l+3 -> half of chunk l/2; channels 128+4l.. -> half of chunk 16 + l/2
    *reinterpret_cast<uint2 *>(tile + tile_chunk_off(rr, l >> 1) + (l & 1) * 8) =
        make_uint2(pack_bf16x2(g0.x, g0.y), pack_bf16x2(g0.z, g0.w));
    *reinterpret_cast<uint2 *>(tile + tile_chunk_off(rr, 16 + (l >> 1)) + (l & 1) * 8) =
        make_uint2(pack_bf16x2(g1.x, g1.y), pack_bf16x2(g1.z, g1.w));
  }
}

// ------------------------------------------------------------------------------------------
// fp32 debug view (parity tests): reference tensor layouts
// ------------------------------------------------------------------------------------------
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
    const float *f00 = sc.features + pix00 * CN_FEAT_C;
    const float *f10 = f00 + (int64_t)W * CN_FEAT_C;
    for (int c = l; c < CN_FEAT_C; c += 32)
      gathered[o * 259 + 3 + c] = lerp2_rn(f00[c], f00[CN_FEAT_C + c], f10[c], f10[CN_FEAT_C + c], pr.ax, pr.ay);
    if (l < 3) {
      const float *i00 = sc.images + pix00 * 3;
      const float *i10 = i00 + (int64_t)W * 3;
      const float tl = __fsub_rn(__fmul_rn(i00[l], 2.f), 1.f), tr = __fsub_rn(__fmul_rn(i00[3 + l], 2.f), 1.f);
      const float bl = __fsub_rn(__fmul_rn(i10[l], 2.f), 1.f), br = __fsub_rn(__fmul_rn(i10[3 + l], 2.f), 1.f);
      gathered[o * 259 + l] = lerp2_rn(tl, tr, bl, br, pr.ax, pr.ay);
    }
  }
}

// ------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <bool kFeatGrad, bool kZGrad>
__global__ void __launch_bounds__(256, 4)
project_gather_bwd_kernel(CnScene sc, const float *__restrict__ ray_o, const float *__restrict__ ray_d,
                          const float *__restrict__ z, int nr, int S, const uint8_t *__restrict__ dx_tiles,
                          float *__restrict__ d_features, float *__restrict__ dz) {
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
  float dz_acc = 0.f;

  for (int v = 0; v < sc.nv; ++v) {
    const int bv = b * sc.nv + v;
    const float *E = sc.extr_inv + (int64_t)bv * 16;
    const float *K = sc.intrinsics + (int64_t)bv * 16;
    const Proj pr = project(E, K, rc.x[0], rc.x[1], rc.x[2], H, W);
    const int64_t pix00 = ((int64_t)bv * H + pr.y0) * W + pr.x0;
    const int64_t row = ((int64_t)bv * nr + r) * S + i;
    const uint8_t *tile = dx_tiles + (row >> 7) * (int64_t)CN_X_TILE_BYTES;
    const uint32_t rr = (uint32_t)(row & 127);
    const uint2 ga = *reinterpret_cast<const uint2 *>(tile + tile_chunk_off(rr, l >> 1) + (l & 1) * 8);
    const uint2 gb = *reinterpret_cast<const uint2 *>(tile + tile_chunk_off(rr, 16 + (l >> 1)) + (l & 1) * 8);
    const float g0[4] = {bf16_lo(ga.x), bf16_hi(ga.x), bf16_lo(ga.y), bf16_hi(ga.y)};
    const float g1[4] = {bf16_lo(gb.x), bf16_hi(gb.x), bf16_lo(gb.y), bf16_hi(gb.y)};

    if (kFeatGrad) {
      float *d00 = d_features + pix00 * CN_FEAT_C;
      float *d10 = d00 + (int64_t)W * CN_FEAT_C;
      const float w00 = (1.f - pr.ax) * (1.f - pr.ay), w01 = pr.ax * (1.f - pr.ay);
      const float w10 = (1.f - pr.ax) * pr.ay, w11 = pr.ax * pr.ay;
      red_add_f4(d00 + 4 * l, w00 * g0[0], w00 * g0[1], w00 * g0[2], w00 * g0[3]);
      red_add_f4(d00 + 128 + 4 * l, w00 * g1[0], w00 * g1[1], w00 * g1[2], w00 * g1[3]);
      red_add_f4(d00 + CN_FEAT_C + 4 * l, w01 * g0[0], w01 * g0[1], w01 * g0[2], w01 * g0[3]);
      red_add_f4(d00 + CN_FEAT_C + 128 + 4 * l, w01 * g1[0], w01 * g1[1], w01 * g1[2], w01 * g1[3]);
      red_add_f4(d10 + 4 * l, w10 * g0[0], w10 * g0[1], w10 * g0[2], w10 * g0[3]);
      red_add_f4(d10 + 128 + 4 * l, w10 * g1[0], w10 * g1[1], w10 * g1[2], w10 * g1[3]);
      red_add_f4(d10 + CN_FEAT_C + 4 * l, w11 * g0[0], w11 * g0[1], w11 * g0[2], w11 * g0[3]);
      red_add_f4(d10 + CN_FEAT_C + 128 + 4 * l, w11 * g1[0], w11 * g1[1], w11 * g1[2], w11 * g1[3]);
    }

    if (kZGrad) {
      // d out / d ax = (1-ay)(TR-TL) + ay(BR-BL);  d out / d ay = (1-ax)(BL-TL) + ax(BR-TR)
      const float *f00 = sc.features + pix00 * CN_FEAT_C;
      const float *f10 = f00 + (int64_t)W * CN_FEAT_C;
      float gax = 0.f, gay = 0.f;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int off = h * 128 + 4 * l;
        const float4 tl = ldg_nc_f4(f00 + off), tr = ldg_nc_f4(f00 + CN_FEAT_C + off);
        const float4 bl = ldg_nc_f4(f10 + off), br = ldg_nc_f4(f10 + CN_FEAT_C + off);
        const float *g = h ? g1 : g0;
        const float tlv[4] = {tl.x, tl.y, tl.z, tl.w}, trv[4] = {tr.x, tr.y, tr.z, tr.w};
        const float blv[4] = {bl.x, bl.y, bl.z, bl.w}, brv[4] = {br.x, br.y, br.z, br.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          gax += g[c] * ((1.f - pr.ay) * (trv[c] - tlv[c]) + pr.ay * (brv[c] - blv[c]));
          gay += g[c] * ((1.f - pr.ax) * (blv[c] - tlv[c]) + pr.ax * (brv[c] - trv[c]));
        }
      }
      float dp[3] = {0.f, 0.f, 0.f};
      if (l < 30) {
        // posenc(position): d/dp_c sin(p_c f_k) = f_k cos, d/dp_c cos(p_c f_k) = -f_k sin
        const int c = l / 10, k = l % 10;
        const float fk = pe_freq(k);
        const uint32_t gp = *reinterpret_cast<const uint32_t *>(tile + tile_off(rr, 256 + 2 * l));
        float s, co;
        sincosf(__fmul_rn(pr.p[c], fk), &s, &co);
        dp[c] = fk * (bf16_lo(gp) * co - bf16_hi(gp) * s);
      } else if (l == 31) {
        const uint4 gi = *reinterpret_cast<const uint4 *>(tile + tile_chunk_off(rr, 47));
        const float gimg[3] = {bf16_lo(gi.x), bf16_hi(gi.x), bf16_lo(gi.y)};
        const float *i00 = sc.images + pix00 * 3;
        const float *i10 = i00 + (int64_t)W * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const float tl = 2.f * i00[c], tr = 2.f * i00[3 + c], bl = 2.f * i10[c], br = 2.f * i10[3 + c];
          gax += gimg[c] * ((1.f - pr.ay) * (tr - tl) + pr.ay * (br - bl));
          gay += gimg[c] * ((1.f - pr.ax) * (bl - tl) + pr.ax * (br - tr));
        }
      }
      gax = warp_sum_f(gax);
      gay = warp_sum_f(gay);
      dp[0] = warp_sum_f(dp[0]);
      dp[1] = warp_sum_f(dp[1]);
      dp[2] = warp_sum_f(dp[2]);
      // alpha clamp, +-1e6 clip and the max(q.z, 1e-8) branch gate the coordinate gradient
      float gpx = (pr.tx >= 0.f && pr.tx <= 1.f && pr.rx >= -1e6f && pr.rx <= 1e6f) ? gax : 0.f;
      float gpy = (pr.ty >= 0.f && pr.ty <= 1.f && pr.ry >= -1e6f && pr.ry <= 1e6f) ? gay : 0.f;
      const float dq0 = gpx / pr.den, dq1 = gpy / pr.den;
      const float dq2 = (pr.q[2] >= 1e-8f) ? -(gpx * pr.rx + gpy * pr.ry) / pr.den : 0.f;
#pragma unroll
      for (int j = 0; j < 3; ++j) dp[j] += K[0 + j] * dq0 + K[4 + j] * dq1 + K[8 + j] * dq2;
      // p = E [o + z d; 1]  ->  dp_i/dz = sum_j E_ij d_j
#pragma unroll
      for (int ii = 0; ii < 3; ++ii)
        dz_acc += dp[ii] * (E[4 * ii + 0] * rc.d[0] + E[4 * ii + 1] * rc.d[1] + E[4 * ii + 2] * rc.d[2]);
    }
  }
  if (kZGrad && l == 0) dz[n] += dz_acc;
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

static int check_scene(const CnScene *s, const char *who) {
  CN_REQUIRE(s, "%s: null scene", who);
  CN_REQUIRE(s->images && s->features && s->intrinsics && s->extr_inv, "%s: null scene pointer", who);
  CN_REQUIRE(s->b > 0 && s->nv > 0 && s->H >= 2 && s->W >= 2, "%s: bad scene dims b=%d nv=%d H=%d W=%d", who, s->b,
             s->nv, s->H, s->W);
  CN_REQUIRE(((uintptr_t)s->features & 15) == 0, "%s: features must be 16-byte aligned", who);
  return 0;
}

}  // namespace cn

extern "C" int cn_project_gather_fwd(const CnScene *scene, const float *ray_o, const float *ray_d, const float *z,
                                     int32_t nr, int32_t S, void *x_tiles, void *stream) {
  using namespace cn;
  if (int e = check_scene(scene, "cn_project_gather_fwd")) return e;
  CN_REQUIRE(ray_o && ray_d && z && x_tiles, "cn_project_gather_fwd: null pointer");
  CN_REQUIRE(nr > 0 && S > 0 && ((int64_t)nr * S) % CN_TILE_ROWS == 0,
             "cn_project_gather_fwd: nr*S must be a positive multiple of %d (nr=%d S=%d)", CN_TILE_ROWS, nr, S);
  const int64_t n = (int64_t)scene->b * nr * S;
  project_gather_fwd_kernel<<<(unsigned)((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(*scene, ray_o, ray_d, z, nr, S,
                                                                                      (uint8_t *)x_tiles);
  return check_launch("cn_project_gather_fwd");
}

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

extern "C" int cn_project_gather_bwd(const CnScene *scene, const float *ray_o, const float *ray_d, const float *z,
                                     int32_t nr, int32_t S, const void *dx_tiles, float *d_features, float *dz,
                                     void *stream) {
  using namespace cn;
  if (int e = check_scene(scene, "cn_project_gather_bwd")) return e;
  CN_REQUIRE(ray_o && ray_d && z && dx_tiles, "cn_project_gather_bwd: null pointer");
  CN_REQUIRE(nr > 0 && S > 0 && ((int64_t)nr * S) % CN_TILE_ROWS == 0,
             "cn_project_gather_bwd: nr*S must be a positive multiple of %d (nr=%d S=%d)", CN_TILE_ROWS, nr, S);
  CN_REQUIRE(!d_features || ((uintptr_t)d_features & 15) == 0, "cn_project_gather_bwd: d_features alignment");
  if (!d_features && !dz) return 0;
  const int64_t n = (int64_t)scene->b * nr * S;
  const unsigned grid = (unsigned)((n + 7) / 8);
  cudaStream_t st = (cudaStream_t)stream;
  const uint8_t *dx = (const uint8_t *)dx_tiles;
  if (d_features && dz)
    project_gather_bwd_kernel<true, true><<<grid, 256, 0, st>>>(*scene, ray_o, ray_d, z, nr, S, dx, d_features, dz);
  else if (d_features)
    project_gather_bwd_kernel<true, false><<<grid, 256, 0, st>>>(*scene, ray_o, ray_d, z, nr, S, dx, d_features, dz);
  else
    project_gather_bwd_kernel<false, true><<<grid, 256, 0, st>>>(*scene, ray_o, ray_d, z, nr, S, dx, d_features, dz);
  return check_launch("cn_project_gather_bwd");
}
