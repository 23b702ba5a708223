// cn_mlp_block_bwd / cn_mlp_readout_bwd / cn_mlp_mean_bwd: the stand-alone kernels of the MLP backward chain.
// The block backward itself is the role body of mlp_bwd_block.cuh (shared with the chained stage kernel).
#include "mlp_bwd_block.cuh"

namespace cn {

template <bool kWgrad>
__global__ void __launch_bounds__(kBwdThreads, 1) mlp_block_bwd_kernel(BlockBwdArgs A) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) BwdBars bars;
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float s_ba[128];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t warp = threadIdx.x >> 5;
  if (threadIdx.x < 128) s_ba[threadIdx.x] = A.bias_a[threadIdx.x];
  if (threadIdx.x == 0) {
    block_bwd_init_barriers(bars);
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
  block_bwd_role<kWgrad>(A, bars, s_ba, smem, tmem, (int)blockIdx.x, (int)gridDim.x);
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------
// SIMT helpers of the backward chain
// ------------------------------------------------------------------------------------------

// readout backward (layers.py:392-397): g_e = (d_y Wr^T) * [relu(e) > 0]; dWr += relu(e)^T d_y; dbr += sum d_y
// Persistent CTAs of 256 threads; the relu(e) tile of the NEXT tile streams into shared memory (bulk TMA, double
// buffered) while this one is processed; g_e leaves through a staged bulk store.
//   phase A: thread = (row, column half): 8 x 16-byte chunks, 4 FMAs per element, packed bf16x2 mask
//   phase B: thread = (column pair, row half): dWr accumulated in registers over all tiles of the CTA
constexpr int kRoThreads = 256;
constexpr uint32_t kRoSmemBytes = 3 * CN_ACT_TILE_BYTES + 1024;

__global__ void __launch_bounds__(kRoThreads, 2)
readout_bwd_kernel(const float *__restrict__ wr, const float *__restrict__ d_y, const uint8_t *__restrict__ acts,
                   int64_t act_stride_bytes, int64_t n_tiles, uint8_t *__restrict__ g_tiles,
                   float *__restrict__ g_wr, float *__restrict__ g_br) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(16) float s_wr[512];
  __shared__ __align__(16) float s_dy[128 * 4];
  __shared__ __align__(8) uint64_t full[2];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t s_base = smem_u32(smem);
  const int tid = threadIdx.x;
  uint8_t *const s_out = smem + 2 * CN_ACT_TILE_BYTES;
  for (int i = tid; i < 512; i += kRoThreads) s_wr[i] = wr[i];
  if (tid == 0) {
    mbar_init(smem_u32(&full[0]), 1);
    mbar_init(smem_u32(&full[1]), 1);
    fence_barrier_init();
  }
  __syncthreads();
  const int64_t n_mine = n_tiles > blockIdx.x ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  auto load = [&](int64_t it) {
    const int64_t tile = blockIdx.x + it * gridDim.x;
    const uint32_t b = (uint32_t)it & 1u;
    mbar_expect_tx(smem_u32(&full[b]), CN_ACT_TILE_BYTES);
    bulk_g2s(s_base + b * CN_ACT_TILE_BYTES, acts + tile * act_stride_bytes, CN_ACT_TILE_BYTES, smem_u32(&full[b]));
  };
  if (tid == 0 && n_mine > 0) load(0);
  // phase A role: row, column half; phase B role: column pair, row half
  const uint32_t row = tid & 127u, half = tid >> 7;
  const uint32_t kp = tid & 63u, rq = tid >> 6;     // columns 2kp, 2kp+1 ; rows [32 rq, 32 rq + 32)
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // dWr[2kp][0..3], dWr[2kp+1][0..3]
  float accb = 0.f;
  for (int64_t it = 0; it < n_mine; ++it) {
    const int64_t tile = blockIdx.x + it * gridDim.x;
    const uint32_t b = (uint32_t)it & 1u;
    const uint8_t *s_a = smem + b * CN_ACT_TILE_BYTES;
    if (tid < 128) reinterpret_cast<float4 *>(s_dy)[tid] = reinterpret_cast<const float4 *>(d_y)[tile * 128 + tid];
    if (tid == 0) {
      bulk_wait_read0();                             // the previous g_e store has left s_out
      if (it + 1 < n_mine) load(it + 1);             // (its buffer was released by the barrier that ended phase B)
    }
    __syncthreads();
    mbar_wait(smem_u32(&full[b]), ((uint32_t)it >> 1) & 1u);
    // ---- phase A
    {
      const float4 dy = reinterpret_cast<const float4 *>(s_dy)[row];
#pragma unroll 2
      for (int c = 0; c < 8; ++c) {
        const uint32_t off = tile_chunk_off(row, half * 8 + c);
        const uint4 a = *reinterpret_cast<const uint4 *>(s_a + off);
        const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
        uint32_t ow[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int k0 = (half * 8 + c) * 8 + 2 * i;
          const float4 w0 = *reinterpret_cast<const float4 *>(s_wr + k0 * 4);
          const float4 w1 = *reinterpret_cast<const float4 *>(s_wr + k0 * 4 + 4);
          const float g0 = dy.x * w0.x + dy.y * w0.y + dy.z * w0.z + dy.w * w0.w;
          const float g1 = dy.x * w1.x + dy.y * w1.y + dy.z * w1.z + dy.w * w1.w;
          ow[i] = mul_bf16x2(cvt_bf16x2(g0, g1), gt0_bf16x2(aw[i]));
        }
        *reinterpret_cast<uint4 *>(s_out + off) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
      }
    }
    // ---- phase B: this thread's two columns over its 32 rows
    {
      const uint32_t chunk = kp >> 2, within = (kp & 3u) * 4u, slab = chunk >> 3;
#pragma unroll 8
      for (uint32_t r = rq * 32u; r < rq * 32u + 32u; ++r) {
        const uint32_t v = *reinterpret_cast<const uint32_t *>(s_a + slab * CN_SLAB_BYTES + r * 128u +
                                                               (((chunk & 7u) ^ (r & 7u)) << 4) + within);
        const float4 dy = reinterpret_cast<const float4 *>(s_dy)[r];
        const float a0 = bf16_lo(v), a1 = bf16_hi(v);
        acc[0] += a0 * dy.x; acc[1] += a0 * dy.y; acc[2] += a0 * dy.z; acc[3] += a0 * dy.w;
        acc[4] += a1 * dy.x; acc[5] += a1 * dy.y; acc[6] += a1 * dy.z; acc[7] += a1 * dy.w;
      }
      if (tid < 4)
        for (int r = 0; r < 128; ++r) accb += s_dy[r * 4 + tid];
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
      bulk_s2g(g_tiles + tile * (int64_t)CN_ACT_TILE_BYTES, s_base + 2 * CN_ACT_TILE_BYTES, CN_ACT_TILE_BYTES);
      bulk_commit();
    }
  }
  if (tid == 0) bulk_wait0();
  // the four row quarters of a column pair meet in shared memory: one atomic per dWr element per CTA
  __syncthreads();
  float *red = reinterpret_cast<float *>(smem);
#pragma unroll
  for (int j = 0; j < 8; ++j) red[(rq * 64 + kp) * 8 + j] = acc[j];
  __syncthreads();
  for (int i = tid; i < 512; i += kRoThreads) {
    const int p2 = i >> 3, j = i & 7;
    const float v = red[(0 * 64 + p2) * 8 + j] + red[(1 * 64 + p2) * 8 + j] + red[(2 * 64 + p2) * 8 + j] +
                    red[(3 * 64 + p2) * 8 + j];
    atomicAdd(&g_wr[(2 * p2 + (j >> 2)) * 4 + (j & 3)], v);
  }
  if (tid < 4) atomicAdd(&g_br[tid], accb);
}

// d(mean over views): g_h3[b][v][t] = g_m[b][t] / nv
__global__ void mean_bwd_kernel(const uint4 *__restrict__ gm, int b, int nv, int64_t tiles_per_view,
                                uint4 *__restrict__ gh3) {
  const int64_t per_b = tiles_per_view * (CN_ACT_TILE_BYTES / 16);
  const int64_t total = (int64_t)b * per_b;
  const float inv = 1.f / (float)nv;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t bb = i / per_b, r = i % per_b;
    const uint4 x = gm[i];
    const uint32_t xw[4] = {x.x, x.y, x.z, x.w};
    uint32_t ow[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) ow[k] = pack_bf16x2(bf16_lo(xw[k]) * inv, bf16_hi(xw[k]) * inv);
    const uint4 o = make_uint4(ow[0], ow[1], ow[2], ow[3]);
    for (int v = 0; v < nv; ++v) gh3[(bb * nv + v) * per_b + r] = o;
  }
}

}  // namespace cn

namespace cn {
CN_TRACE_READER(cn_debug_read_trace)
}

extern "C" int cn_mlp_block_bwd(const void *packed, int32_t blk, const void *acts, int32_t act_stride_tiles,
                                int32_t act_index, int64_t n_tiles, const void *gy_tiles, int32_t gy_views,
                                int64_t gy_tiles_per_view, void *gh_tiles, float scale, float *grad, void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && acts && gy_tiles && gh_tiles, "cn_mlp_block_bwd: null pointer");
  CN_REQUIRE(blk >= 0 && blk < 6, "cn_mlp_block_bwd: blk %d", blk);
  CN_REQUIRE(act_stride_tiles >= 1 && act_index >= 0 && act_index < act_stride_tiles,
             "cn_mlp_block_bwd: act_stride_tiles=%d act_index=%d", act_stride_tiles, act_index);
  CN_REQUIRE(n_tiles >= 0, "cn_mlp_block_bwd: n_tiles %lld", (long long)n_tiles);
  CN_REQUIRE(gy_views >= 1 && (gy_views == 1 || (gy_tiles_per_view > 0 && n_tiles % ((int64_t)gy_views * gy_tiles_per_view) == 0 &&
                                                 gy_tiles != gh_tiles)),
             "cn_mlp_block_bwd: gy_views=%d gy_tiles_per_view=%lld (shared g_y needs a separate g_h buffer)", gy_views,
             (long long)gy_tiles_per_view);
  CN_REQUIRE((((uintptr_t)packed | (uintptr_t)acts | (uintptr_t)gy_tiles | (uintptr_t)gh_tiles) & 127) == 0 &&
                 ((uintptr_t)grad & 15) == 0,
             "cn_mlp_block_bwd: alignment");
  if (n_tiles == 0) return 0;
  cudaError_t e = grad ? cudaFuncSetAttribute(mlp_block_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)kBwdSmemBytes)
                       : cudaFuncSetAttribute(mlp_block_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)kBwdSmemBytes);
  if (e != cudaSuccess) {
    set_error("cn_mlp_block_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  const uint8_t *pk = (const uint8_t *)packed;
  const unsigned grid = (unsigned)(n_tiles < num_sms() ? n_tiles : num_sms());
  // grad == NULL: gradient to the block input only (no weight / bias gradients are formed)
  static float dummy_grad;   // block_bwd_args only forms addresses from it
  BlockBwdArgs A = block_bwd_args(pk, blk, (const uint8_t *)acts, act_stride_tiles, act_index, n_tiles, scale,
                                  grad ? grad : &dummy_grad);
  A.gy_tiles = (const uint8_t *)gy_tiles;
  A.gy_views = gy_views;
  A.gy_tpv = gy_tiles_per_view;
  A.gh_tiles = (uint8_t *)gh_tiles;
  if (grad) mlp_block_bwd_kernel<true><<<grid, kBwdThreads, kBwdSmemBytes, (cudaStream_t)stream>>>(A);
  else mlp_block_bwd_kernel<false><<<grid, kBwdThreads, kBwdSmemBytes, (cudaStream_t)stream>>>(A);
  return check_launch("cn_mlp_block_bwd");
}

extern "C" int cn_mlp_readout_bwd(const float *params, const float *d_y, const void *acts, int32_t act_stride_tiles,
                                  int32_t act_index, int64_t n_tiles, void *g_tiles, float *grad, void *stream) {
  using namespace cn;
  CN_REQUIRE(params && d_y && acts && g_tiles && grad, "cn_mlp_readout_bwd: null pointer");
  CN_REQUIRE(act_stride_tiles >= 1 && act_index >= 0 && act_index < act_stride_tiles,
             "cn_mlp_readout_bwd: act_stride_tiles=%d act_index=%d", act_stride_tiles, act_index);
  CN_REQUIRE(n_tiles >= 0, "cn_mlp_readout_bwd: n_tiles %lld", (long long)n_tiles);
  CN_REQUIRE((((uintptr_t)d_y | (uintptr_t)acts | (uintptr_t)g_tiles) & 127) == 0, "cn_mlp_readout_bwd: alignment");
  if (n_tiles == 0) return 0;
  cudaError_t e = cudaFuncSetAttribute(readout_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRoSmemBytes);
  if (e != cudaSuccess) {
    set_error("cn_mlp_readout_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  const unsigned grid = (unsigned)(n_tiles < 2 * num_sms() ? n_tiles : 2 * num_sms());
  readout_bwd_kernel<<<grid, kRoThreads, kRoSmemBytes, (cudaStream_t)stream>>>(
      params + kParamWr, d_y, (const uint8_t *)acts + (int64_t)act_index * CN_ACT_TILE_BYTES,
      (int64_t)act_stride_tiles * CN_ACT_TILE_BYTES, n_tiles, (uint8_t *)g_tiles, grad + kParamWr, grad + kParamBr);
  return check_launch("cn_mlp_readout_bwd");
}

extern "C" int cn_mlp_mean_bwd(const void *gm_tiles, int32_t b, int32_t nv, int64_t tiles_per_view, void *gh3_tiles,
                               void *stream) {
  using namespace cn;
  CN_REQUIRE(gm_tiles && gh3_tiles, "cn_mlp_mean_bwd: null pointer");
  CN_REQUIRE(b > 0 && nv > 0 && tiles_per_view >= 0, "cn_mlp_mean_bwd: b=%d nv=%d tiles_per_view=%lld", b, nv,
             (long long)tiles_per_view);
  if (tiles_per_view == 0) return 0;
  mean_bwd_kernel<<<num_sms() * 8, 256, 0, (cudaStream_t)stream>>>((const uint4 *)gm_tiles, b, nv, tiles_per_view,
                                                                 (uint4 *)gh3_tiles);
  return check_launch("cn_mlp_mean_bwd");
}
