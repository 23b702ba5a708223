// Full-image bookkeeping of render_view (model_v0.py:243-281) on the device: rays of every pixel of the target
// view (get_rays / get_specific_rays, nerf_utils.py:15-35) and the uint8 conversion of the rendered image
// (model_v0.py:272-280).  The reference generates the rays with numpy on the host, walks the image in 600 chunks
// of 512 rays and copies every chunk's floats back; here the rays never exist on the host and only the two uint8
// images leave the device.
#include "common.cuh"

namespace cn {

struct RayGen {
  double m[9];   // R . K^-1 (row-major), computed on the host in float64 like the reference's numpy expression
  float o[3];    // camera centre pose[:3, 3]
};

// pixel p (row-major, u = column, v = row, nerf_utils.py:16-19): d = m . [u, v, 1] / |.|, in the arithmetic numpy
// uses for `(pose[:3,:3] @ inv(K) @ pixels)`: float64 when the pose is float64 (numpy promotes the float32 pixel
// grid), float32 when pose and K are both float32; rounded to fp32 once at the end
template <typename T>
__global__ void __launch_bounds__(256) generate_rays_kernel(RayGen g, int W, int64_t first, int64_t n,
                                                            float *__restrict__ ray_o, float *__restrict__ ray_d) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const int64_t p = first + i;
  const T u = (T)(float)(p % W), v = (T)(float)(p / W);
  T d[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) d[r] = (T)g.m[3 * r] * u + (T)g.m[3 * r + 1] * v + (T)g.m[3 * r + 2];
  const T nrm = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    ray_d[i * 3 + r] = (float)(d[r] / nrm);
    ray_o[i * 3 + r] = g.o[r];
  }
}

// order-preserving map float -> uint32 (for atomicMin / atomicMax on floats of either sign)
__device__ __forceinline__ uint32_t f2ord(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o) {
  return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

__global__ void __launch_bounds__(256) depth_minmax_kernel(const float *__restrict__ depth, int64_t n,
                                                           uint32_t *__restrict__ mm) {
  uint32_t lo = 0xffffffffu, hi = 0u;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    const uint32_t o = f2ord(depth[i]);
    lo = min(lo, o);
    hi = max(hi, o);
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, s));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, s));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(mm, lo);
    atomicMax(mm + 1, hi);
  }
}

// all_rgbs * 255 -> clip [0, 255] -> astype(uint8) (truncation); (depth - min) / (max - min) * 255 -> astype(uint8),
// every step a separately rounded fp32 operation as in numpy
__global__ void __launch_bounds__(256) quantize_kernel(const float *__restrict__ rgb, const float *__restrict__ depth,
                                                       int64_t n, const uint32_t *__restrict__ mm,
                                                       uint8_t *__restrict__ rgb8, uint8_t *__restrict__ depth8) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float v = fminf(fmaxf(__fmul_rn(rgb[i * 3 + c], 255.f), 0.f), 255.f);
    rgb8[i * 3 + c] = (uint8_t)(int)v;
  }
  const float lo = ord2f(mm[0]), hi = ord2f(mm[1]);
  const float nd = __fdiv_rn(__fsub_rn(depth[i], lo), __fsub_rn(hi, lo));
  depth8[i] = (uint8_t)(int)__fmul_rn(nd, 255.f);
}

// images of MVNeRFDataGenerator.get_input (mvnerf.py:37-38): np.array(colors) / 255.0 (float64) cast to float32 --
// evaluated on the device from the dataset's uint8 pixels (first 3 of src_channels channels), bit-identical to numpy
__global__ void __launch_bounds__(256) images_u8_kernel(const uint8_t *__restrict__ src, int64_t n_pixels,
                                                        int src_channels, float *__restrict__ dst) {
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n_pixels * 3; i += (int64_t)gridDim.x * 256) {
    const int64_t p = i / 3;
    const int c = (int)(i - p * 3);
    dst[i] = (float)((double)src[p * src_channels + c] / 255.0);
  }
}

}  // namespace cn

extern "C" int cn_images_u8_to_f32(const uint8_t *src, int64_t n_pixels, int32_t src_channels, float *dst,
                                   void *stream) {
  using namespace cn;
  CN_REQUIRE(src && dst, "cn_images_u8_to_f32: null pointer");
  CN_REQUIRE(n_pixels >= 0 && src_channels >= 3, "cn_images_u8_to_f32: n_pixels=%lld src_channels=%d",
             (long long)n_pixels, src_channels);
  if (n_pixels == 0) return 0;
  const int64_t want = (n_pixels * 3 + 255) / 256;
  const unsigned blocks = (unsigned)(want < 16 * num_sms() ? want : 16 * num_sms());
  images_u8_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(src, n_pixels, src_channels, dst);
  return check_launch("cn_images_u8_to_f32");
}

extern "C" int cn_generate_rays(const double *rk_inv_host, const float *origin_host, int32_t single_precision, int32_t W,
                                int64_t first_pixel, int64_t n, float *ray_o, float *ray_d, void *stream) {
  using namespace cn;
  CN_REQUIRE(rk_inv_host && origin_host && ray_o && ray_d, "cn_generate_rays: null pointer");
  CN_REQUIRE(W > 0 && first_pixel >= 0 && n >= 0, "cn_generate_rays: W=%d first=%lld n=%lld", W, (long long)first_pixel,
             (long long)n);
  if (n == 0) return 0;
  RayGen g;
  for (int i = 0; i < 9; ++i) g.m[i] = rk_inv_host[i];
  for (int i = 0; i < 3; ++i) g.o[i] = origin_host[i];
  const unsigned blocks = (unsigned)((n + 255) / 256);
  if (single_precision)
    generate_rays_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>(g, W, first_pixel, n, ray_o, ray_d);
  else
    generate_rays_kernel<double><<<blocks, 256, 0, (cudaStream_t)stream>>>(g, W, first_pixel, n, ray_o, ray_d);
  return check_launch("cn_generate_rays");
}

extern "C" int cn_quantize_render(const float *rgb, const float *depth, int64_t n, void *minmax_ws, uint8_t *rgb8,
                                  uint8_t *depth8, void *stream) {
  using namespace cn;
  CN_REQUIRE(rgb && depth && minmax_ws && rgb8 && depth8, "cn_quantize_render: null pointer");
  CN_REQUIRE(n > 0, "cn_quantize_render: n=%lld", (long long)n);
  CN_REQUIRE(((uintptr_t)minmax_ws & 7) == 0, "cn_quantize_render: minmax_ws must be an 8-byte aligned 8-byte buffer");
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(minmax_ws, 0xff, 4, st);                       // running min (ordered uint)
  if (e == cudaSuccess) e = cudaMemsetAsync((uint8_t *)minmax_ws + 4, 0, 4, st);   // running max
  if (e != cudaSuccess) {
    set_error("cn_quantize_render: cudaMemsetAsync: %s", cudaGetErrorString(e));
    return (int)e;
  }
  const unsigned blocks = (unsigned)((n + 255) / 256);
  depth_minmax_kernel<<<blocks < 1184u ? blocks : 1184u, 256, 0, st>>>(depth, n, (uint32_t *)minmax_ws);
  quantize_kernel<<<blocks, 256, 0, st>>>(rgb, depth, n, (const uint32_t *)minmax_ws, rgb8, depth8);
  return check_launch("cn_quantize_render");
}
