// Device helpers shared by the projection / gather kernels: explicitly rounded projection of a
// world point into a source view (compute_pixel_in_image_mv, nerf_utils.py:64-81), the edge-clamped
// bilinear footprint of tensorflow_addons.image.interpolate_bilinear, and the posenc frequencies
// (nerf_utils.py:108-126).
#pragma once
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


__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// sin / cos of a posenc argument (|a| up to a few 1e3 rad): Cody-Waite reduction by multiples of pi
// to [-pi/2, pi/2], then the SFU approximations (abs err ~5e-7 there; the results are stored as
// bf16).  Huge or non-finite arguments take the libm path.
__device__ __forceinline__ void sincos_posenc(float a, float &s, float &c) {
  if (!(fabsf(a) < 5.0e4f)) {
    sincosf(a, &s, &c);
    return;
  }
  const float n = rintf(a * 0.318309886183790672f);
  float r = fmaf(n, -3.140625f, a);
  r = fmaf(n, -9.67502593994140625e-4f, r);
  r = fmaf(n, -1.509957990978376e-7f, r);
  s = __sinf(r);
  c = __cosf(r);
  if (((int)n) & 1) {
    s = -s;
    c = -c;
  }
}

}  // namespace cn
