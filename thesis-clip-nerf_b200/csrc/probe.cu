// cn_probe_umma: one 128 x 128 accumulator tile computed by n_mma tcgen05.mma instructions
// from caller-supplied shared-memory operand images and descriptor fields.  tests/ uses it to
// pin the descriptor encoding (K-major and MN-major, 128-byte swizzle) that the MLP kernels
// rely on, against a CPU matmul.
#include <cstdio>

#include "common.cuh"

namespace cn {

constexpr uint32_t kProbeOperandBytes = 98304;

__global__ void __launch_bounds__(128, 1)
probe_umma_kernel(const uint8_t *__restrict__ a_img, uint32_t a_bytes, const uint8_t *__restrict__ b_img,
                  uint32_t b_bytes, int n_mma, uint32_t idesc, uint32_t a_lbo, uint32_t a_sbo, uint32_t a_kstep,
                  uint32_t b_lbo, uint32_t b_sbo, uint32_t b_kstep, float *__restrict__ d_out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_base_slot;
  const uint32_t sA = (smem_u32(smem_raw) + 1023u) & ~1023u;  // swizzle atoms need 1024-byte alignment
  const uint32_t sB = sA + kProbeOperandBytes;
  const uint32_t bar_load = smem_u32(&bars[0]);
  const uint32_t bar_mma = smem_u32(&bars[1]);
  const uint32_t warp = threadIdx.x >> 5;

  if (threadIdx.x == 0) {
    mbar_init(bar_load, 1);
    mbar_init(bar_mma, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&tmem_base_slot), 128);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_slot;

  if (threadIdx.x == 0) {
    mbar_expect_tx(bar_load, a_bytes + b_bytes);
    for (uint32_t off = 0; off < a_bytes; off += CN_SLAB_BYTES)
      bulk_g2s(sA + off, a_img + off, min(a_bytes - off, (uint32_t)CN_SLAB_BYTES), bar_load);
    for (uint32_t off = 0; off < b_bytes; off += CN_SLAB_BYTES)
      bulk_g2s(sB + off, b_img + off, min(b_bytes - off, (uint32_t)CN_SLAB_BYTES), bar_load);
    mbar_wait(bar_load, 0);
    tc_fence_after();
    for (int j = 0; j < n_mma; ++j) {
      const uint64_t da = umma_desc(sA + j * a_kstep, a_lbo, a_sbo);
      const uint64_t db = umma_desc(sB + j * b_kstep, b_lbo, b_sbo);
      umma_f16(tmem, da, db, idesc, j > 0 ? 1u : 0u);
    }
    umma_commit(bar_mma);
  }
  __syncwarp();
  mbar_wait(bar_mma, 0);
  tc_fence_after();

  const uint32_t row = threadIdx.x;  // warp w owns TMEM lanes [32w, 32w+32)
  for (int c = 0; c < 4; ++c) {
    uint32_t r[32];
    tmem_ld32(tmem + ((warp * 32u) << 16) + c * 32u, r);
    tmem_wait_ld();
#pragma unroll
    for (int j = 0; j < 32; ++j) d_out[row * 128 + c * 32 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

// ---- tile reduce probe: 81 KiB tiles added to a [n_img, H, W, 256] fp32 map
constexpr int kProbeTT = 9;
__global__ void __launch_bounds__(256)
probe_tile_reduce_kernel(float *__restrict__ map, int n_img, int H, int W, int iters, int mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  float *tile = reinterpret_cast<float *>(smem);
  for (int i = threadIdx.x; i < kProbeTT * kProbeTT * 256; i += blockDim.x) tile[i] = 1.f;
  fence_proxy_async_smem();
  __syncthreads();
  const int tiles_x = (W - 1) / 8, tiles_y = (H - 1) / 8;
  uint32_t rng = 0x9e3779b9u * (blockIdx.x + 1);
  for (int it = 0; it < iters; ++it) {
    rng = rng * 1664525u + 1013904223u;
    const int tx = (rng >> 8) % tiles_x;
    rng = rng * 1664525u + 1013904223u;
    const int ty = (rng >> 8) % tiles_y;
    rng = rng * 1664525u + 1013904223u;
    const int img = (rng >> 8) % n_img;
    float *dst = map + (((int64_t)img * H + ty * 8) * W + tx * 8) * 256;
    if (mode == 0) {
      if (threadIdx.x == 0) {
        for (int yy = 0; yy < kProbeTT; ++yy)
          bulk_reduce_add_f32(dst + (int64_t)yy * W * 256, smem_u32(tile) + yy * kProbeTT * 1024, kProbeTT * 1024);
        bulk_commit();
        bulk_wait_read0();
      }
      __syncthreads();
    } else {
      // 81 texels x 64 float4 per texel, one float4 per thread per step
      for (int i = threadIdx.x; i < kProbeTT * kProbeTT * 64; i += blockDim.x) {
        const int tex = i >> 6, c4 = i & 63;
        const int y = tex / kProbeTT, x = tex % kProbeTT;
        red_add_f4(dst + ((int64_t)y * W + x) * 256 + c4 * 4, 1.f, 1.f, 1.f, 1.f);
      }
    }
  }
  if (mode == 0 && threadIdx.x == 0) bulk_wait0();
}

}  // namespace cn

extern "C" int cn_probe_tile_reduce(float *map, int32_t n_img, int32_t H, int32_t W, int32_t iters, int32_t mode,
                                    int32_t ctas_per_sm, void *stream) {
  using namespace cn;
  CN_REQUIRE(map && n_img > 0 && H >= 17 && W >= 17 && iters > 0 && ctas_per_sm > 0 && ctas_per_sm <= 2,
             "cn_probe_tile_reduce: bad arguments");
  const size_t smem = kProbeTT * kProbeTT * 1024 + 1024;
  cudaError_t e = cudaFuncSetAttribute(probe_tile_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) {
    set_error("cn_probe_tile_reduce: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  probe_tile_reduce_kernel<<<num_sms() * ctas_per_sm, 256, smem, (cudaStream_t)stream>>>(map, n_img, H, W, iters, mode);
  return check_launch("cn_probe_tile_reduce");
}

extern "C" int cn_probe_umma(const void *a_img, int32_t a_bytes, const void *b_img, int32_t b_bytes, int32_t n_mma,
                             int32_t a_major_mn, int32_t b_major_mn, int32_t a_lbo, int32_t a_sbo, int32_t a_kstep,
                             int32_t b_lbo, int32_t b_sbo, int32_t b_kstep, float *d_out, void *stream) {
  using namespace cn;
  CN_REQUIRE(a_img && b_img && d_out, "cn_probe_umma: null pointer");
  CN_REQUIRE(a_bytes > 0 && a_bytes <= (int)kProbeOperandBytes && a_bytes % 16 == 0, "cn_probe_umma: a_bytes %d",
             a_bytes);
  CN_REQUIRE(b_bytes > 0 && b_bytes <= (int)kProbeOperandBytes && b_bytes % 16 == 0, "cn_probe_umma: b_bytes %d",
             b_bytes);
  CN_REQUIRE(n_mma >= 1 && n_mma <= 64, "cn_probe_umma: n_mma %d", n_mma);
  const size_t smem = 2 * kProbeOperandBytes + 1024;
  cudaError_t e = cudaFuncSetAttribute(probe_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) {
    set_error("cn_probe_umma: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  const uint32_t idesc = umma_idesc(128, 128, a_major_mn ? 1u : 0u, b_major_mn ? 1u : 0u);
  probe_umma_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(
      (const uint8_t *)a_img, (uint32_t)a_bytes, (const uint8_t *)b_img, (uint32_t)b_bytes, n_mma, idesc,
      (uint32_t)a_lbo, (uint32_t)a_sbo, (uint32_t)a_kstep, (uint32_t)b_lbo, (uint32_t)b_sbo, (uint32_t)b_kstep, d_out);
  return check_launch("cn_probe_umma");
}
