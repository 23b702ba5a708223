// Alpha compositing with a transmittance scan (model_v0.py:89-100, nerf_utils.py:129-140),
// forward and backward.  One warp per ray, S/32 consecutive samples per lane; the exclusive
// product scan over samples is a per-lane serial product + a warp shuffle scan.  The backward
// recomputes alpha / T / weights (nothing but z and the readout outputs is read) and uses a
// reverse (suffix) shuffle scan for the "everything behind this sample" term.
#include "common.cuh"

namespace cn {

template <int SPL>
struct RaySamples {
  float z[SPL], dlt[SPL], sig[SPL], c0[SPL], c1[SPL], c2[SPL];
  float alpha[SPL], e[SPL], f[SPL], T[SPL], w[SPL];
};

template <int SPL>
__device__ __forceinline__ void load_and_scan(const float *__restrict__ z, const float *__restrict__ rgbs,
                                              int64_t ray, RaySamples<SPL> &s) {
  constexpr int S = SPL * 32;
  const int l = lane_id();
  const float *zr = z + ray * S + l * SPL;
  const float4 *cr = reinterpret_cast<const float4 *>(rgbs) + ray * S + l * SPL;
#pragma unroll
  for (int k = 0; k < SPL; ++k) {
    s.z[k] = zr[k];
    const float4 v = cr[k];
    s.c0[k] = v.x; s.c1[k] = v.y; s.c2[k] = v.z; s.sig[k] = v.w;
  }
  // next z: own next sample, or the first sample of the next lane
  const float z_next_lane = __shfl_down_sync(0xffffffffu, s.z[0], 1);
#pragma unroll
  for (int k = 0; k < SPL; ++k) {
    const float zn = (k + 1 < SPL) ? s.z[(k + 1) % SPL] : z_next_lane;
    s.dlt[k] = zn - s.z[k];
  }
  // last interval repeats the previous one (dists[..., -1:] concatenated)
  {
    const float prev_last = (SPL > 1) ? s.dlt[(SPL > 1) ? SPL - 2 : 0] : __shfl_up_sync(0xffffffffu, s.dlt[0], 1);
    if (l == 31) s.dlt[SPL - 1] = prev_last;
  }
  float prod = 1.f;
#pragma unroll
  for (int k = 0; k < SPL; ++k) {
    s.e[k] = expf(-s.dlt[k] * fmaxf(s.sig[k], 0.f));
    s.alpha[k] = 1.f - s.e[k];
    s.f[k] = (1.f - s.alpha[k]) + 1e-10f;
    s.T[k] = prod;  // exclusive product inside the lane
    prod *= s.f[k];
  }
  // exclusive product scan across lanes
  float incl = prod;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float v = __shfl_up_sync(0xffffffffu, incl, o);
    if (l >= o) incl *= v;
  }
  float excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (l == 0) excl = 1.f;
#pragma unroll
  for (int k = 0; k < SPL; ++k) {
    s.T[k] *= excl;
    s.w[k] = s.alpha[k] * s.T[k];
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int SPL>
__global__ void __launch_bounds__(128)
composite_fwd_kernel(const float *__restrict__ z, const float *__restrict__ rgbs, int64_t n_rays,
                     float *__restrict__ rgb, float *__restrict__ depth, float *__restrict__ weights) {
  constexpr int S = SPL * 32;
  const int64_t ray = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (ray >= n_rays) return;
  const int l = lane_id();
  RaySamples<SPL> s;
  load_and_scan<SPL>(z, rgbs, ray, s);
  float r = 0.f, g = 0.f, b = 0.f, d = 0.f;
#pragma unroll
  for (int k = 0; k < SPL; ++k) {
    r += s.w[k] * s.c0[k];
    g += s.w[k] * s.c1[k];
    b += s.w[k] * s.c2[k];
    d += s.w[k] * s.z[k];
    weights[ray * S + l * SPL + k] = s.w[k];
  }
  r = warp_sum(r); g = warp_sum(g); b = warp_sum(b); d = warp_sum(d);
  if (l == 0) {
    rgb[ray * 3 + 0] = r; rgb[ray * 3 + 1] = g; rgb[ray * 3 + 2] = b;
    depth[ray] = d;
  }
}

template <int SPL>
__global__ void __launch_bounds__(128)
composite_bwd_kernel(const float *__restrict__ z, const float *__restrict__ rgbs, int64_t n_rays,
                     const float *__restrict__ d_rgb, const float *__restrict__ d_depth,
                     const float *__restrict__ d_weights, float *__restrict__ d_y, float *__restrict__ dz) {
  constexpr int S = SPL * 32;
  const int64_t ray = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (ray >= n_rays) return;
  const int l = lane_id();
  RaySamples<SPL> s;
  load_and_scan<SPL>(z, rgbs, ray, s);
  const float gr = d_rgb[ray * 3 + 0], gg = d_rgb[ray * 3 + 1], gb = d_rgb[ray * 3 + 2];
  const float gd = d_depth ? d_depth[ray] : 0.f;
  float gw[SPL], suf[SPL];
  float run = 0.f;
  // suffix sums of g_w * w strictly behind each sample: serial inside the lane, shuffle scan across lanes
#pragma unroll
  for (int k = SPL - 1; k >= 0; --k) {
    gw[k] = gr * s.c0[k] + gg * s.c1[k] + gb * s.c2[k] + gd * s.z[k];
    if (d_weights) gw[k] += d_weights[ray * S + l * SPL + k];
    suf[k] = run;
    run += gw[k] * s.w[k];
  }
  float incl = run;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float v = __shfl_down_sync(0xffffffffu, incl, o);
    if (l + o < 32) incl += v;
  }
  float behind = __shfl_down_sync(0xffffffffu, incl, 1);
  if (l == 31) behind = 0.f;
  float ddl[SPL];
#pragma unroll
  for (int k = 0; k < SPL; ++k) {
    const float S_i = suf[k] + behind;
    const float dalpha = gw[k] * s.T[k] - S_i / s.f[k];
    const float relu_s = fmaxf(s.sig[k], 0.f);
    const float dsig = (s.sig[k] > 0.f) ? dalpha * s.dlt[k] * s.e[k] : 0.f;
    ddl[k] = dalpha * relu_s * s.e[k];
    float4 o;
    o.x = s.w[k] * gr * s.c0[k] * (1.f - s.c0[k]);  // sigmoid'
    o.y = s.w[k] * gg * s.c1[k] * (1.f - s.c1[k]);
    o.z = s.w[k] * gb * s.c2[k] * (1.f - s.c2[k]);
    o.w = dsig * (1.f - expf(-s.sig[k]));  // softplus' = sigmoid(y) = 1 - exp(-softplus(y))
    reinterpret_cast<float4 *>(d_y)[ray * S + l * SPL + k] = o;
  }
  if (dz) {
    // delta_i = z_{i+1} - z_i (i < S-1), delta_{S-1} = delta_{S-2}
    const float last = __shfl_sync(0xffffffffu, ddl[SPL - 1], 31);
    if (SPL > 1) {
      if (l == 31) { ddl[(SPL > 1) ? SPL - 2 : 0] += last; ddl[SPL - 1] = 0.f; }
    } else {
      if (l == 30) ddl[0] += last;
      if (l == 31) ddl[0] = 0.f;
    }
    float prev_lane_last = __shfl_up_sync(0xffffffffu, ddl[SPL - 1], 1);
    if (l == 0) prev_lane_last = 0.f;
#pragma unroll
    for (int k = 0; k < SPL; ++k) {
      const float prev = (k == 0) ? prev_lane_last : ddl[(k + SPL - 1) % SPL];
      dz[ray * S + l * SPL + k] = gd * s.w[k] + prev - ddl[k];
    }
  }
}

}  // namespace cn

extern "C" int cn_composite_fwd(const float *z, const float *rgbs, int64_t n_rays, int32_t S, float *rgb,
                                float *depth, float *weights, void *stream) {
  using namespace cn;
  CN_REQUIRE(z && rgbs && rgb && depth && weights, "cn_composite_fwd: null pointer");
  CN_REQUIRE(S == 64 || S == 128, "cn_composite_fwd: S must be 64 or 128 (got %d)", S);
  CN_REQUIRE(n_rays >= 0, "cn_composite_fwd: n_rays %lld", (long long)n_rays);
  if (n_rays == 0) return 0;
  const unsigned grid = (unsigned)((n_rays + 3) / 4);
  if (S == 64)
    composite_fwd_kernel<2><<<grid, 128, 0, (cudaStream_t)stream>>>(z, rgbs, n_rays, rgb, depth, weights);
  else
    composite_fwd_kernel<4><<<grid, 128, 0, (cudaStream_t)stream>>>(z, rgbs, n_rays, rgb, depth, weights);
  return check_launch("cn_composite_fwd");
}

extern "C" int cn_composite_bwd(const float *z, const float *rgbs, int64_t n_rays, int32_t S, const float *d_rgb,
                                const float *d_depth, const float *d_weights, float *d_y, float *dz, void *stream) {
  using namespace cn;
  CN_REQUIRE(z && rgbs && d_rgb && d_y, "cn_composite_bwd: null pointer");
  CN_REQUIRE(S == 64 || S == 128, "cn_composite_bwd: S must be 64 or 128 (got %d)", S);
  CN_REQUIRE(n_rays >= 0, "cn_composite_bwd: n_rays %lld", (long long)n_rays);
  if (n_rays == 0) return 0;
  const unsigned grid = (unsigned)((n_rays + 3) / 4);
  if (S == 64)
    composite_bwd_kernel<2><<<grid, 128, 0, (cudaStream_t)stream>>>(z, rgbs, n_rays, d_rgb, d_depth, d_weights, d_y,
                                                                     dz);
  else
    composite_bwd_kernel<4><<<grid, 128, 0, (cudaStream_t)stream>>>(z, rgbs, n_rays, d_rgb, d_depth, d_weights, d_y,
                                                                     dz);
  return check_launch("cn_composite_bwd");
}
