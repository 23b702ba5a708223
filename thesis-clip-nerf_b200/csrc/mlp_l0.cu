// cn_mlp_l0_dgrad / cn_mlp_l0_wgrad: stand-alone kernels of the first Dense's backward (role bodies in mlp_l0.cuh).
#include "mlp_l0.cuh"

namespace cn {

__global__ void __launch_bounds__(kDgThreads, 1) mlp_l0_dgrad_kernel(L0DgradArgs A) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) DgBars bars;
  __shared__ uint32_t tmem_slot;
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    l0_dgrad_init_barriers(bars);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  l0_dgrad_role(A, bars, smem, tmem, (int)blockIdx.x, (int)gridDim.x);
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

__global__ void __launch_bounds__(kWgThreads, 1) mlp_l0_wgrad_kernel(L0WgradArgs A) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) WgBars bars;
  __shared__ uint32_t tmem_slot;
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    l0_wgrad_init_barriers(bars);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  l0_wgrad_role(A, bars, smem, tmem, (int)blockIdx.x, (int)gridDim.x);
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

}  // namespace cn

extern "C" int cn_mlp_l0_dgrad(const void *packed, const void *g0_tiles, int64_t n_tiles, void *dx_tiles,
                               float scale, void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && g0_tiles && dx_tiles, "cn_mlp_l0_dgrad: null pointer");
  CN_REQUIRE(n_tiles >= 0, "cn_mlp_l0_dgrad: n_tiles %lld", (long long)n_tiles);
  CN_REQUIRE((((uintptr_t)packed | (uintptr_t)g0_tiles | (uintptr_t)dx_tiles) & 127) == 0, "cn_mlp_l0_dgrad: alignment");
  if (n_tiles == 0) return 0;
  cudaError_t e =
      cudaFuncSetAttribute(mlp_l0_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDgSmemBytes);
  if (e != cudaSuccess) {
    set_error("cn_mlp_l0_dgrad: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  const unsigned grid = (unsigned)(n_tiles < num_sms() ? n_tiles : num_sms());
  L0DgradArgs A{};
  A.w0d = (const uint8_t *)packed + kPackBwdW0;
  A.g0_tiles = (const uint8_t *)g0_tiles;
  A.n_tiles = n_tiles;
  A.dx_tiles = (uint8_t *)dx_tiles;
  A.scale = scale;
  mlp_l0_dgrad_kernel<<<grid, kDgThreads, kDgSmemBytes, (cudaStream_t)stream>>>(A);
  return check_launch("cn_mlp_l0_dgrad");
}

extern "C" int cn_mlp_l0_wgrad(const void *x_tiles, const void *g0_tiles, int64_t n_tiles, float scale, float *grad,
                               void *stream) {
  using namespace cn;
  CN_REQUIRE(x_tiles && g0_tiles && grad, "cn_mlp_l0_wgrad: null pointer");
  CN_REQUIRE(n_tiles >= 0, "cn_mlp_l0_wgrad: n_tiles %lld", (long long)n_tiles);
  CN_REQUIRE((((uintptr_t)x_tiles | (uintptr_t)g0_tiles) & 127) == 0 && ((uintptr_t)grad & 15) == 0,
             "cn_mlp_l0_wgrad: alignment");
  if (n_tiles == 0) return 0;
  cudaError_t e =
      cudaFuncSetAttribute(mlp_l0_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWgSmemBytes);
  if (e != cudaSuccess) {
    set_error("cn_mlp_l0_wgrad: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  const unsigned grid = (unsigned)(n_tiles < num_sms() ? n_tiles : num_sms());
  L0WgradArgs A{};
  A.x_tiles = (const uint8_t *)x_tiles;
  A.g0_tiles = (const uint8_t *)g0_tiles;
  A.n_tiles = n_tiles;
  A.g_w0 = grad + kParamW0;
  A.g_b0 = grad + kParamB0;
  A.scale = scale;
  mlp_l0_wgrad_kernel<<<grid, kWgThreads, kWgSmemBytes, (cudaStream_t)stream>>>(A);
  return check_launch("cn_mlp_l0_wgrad");
}
