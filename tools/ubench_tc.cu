// Micro-benchmarks that pin the per-SM cost model the MLP kernels are designed against (B200, sm_100a):
//   mode 0: back-to-back tcgen05.mma 128x128x16 (cycles per MMA)
//   mode 1: tcgen05.ld 32x32b.x32 throughput, W warps reading 128x128 fp32 tiles (cycles per tile)
//   mode 2: full epilogue (TMEM -> +bias -> relu -> bf16 -> swizzled st.shared), W warps (cycles per tile)
//   mode 3: mode 2 while one thread keeps the tensor pipe busy (interference)
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_tc tools/ubench_tc.cu
#include <cstdio>
#include <cstdlib>
#include "../thesis-clip-nerf_b200/csrc/common.cuh"

namespace cn {
void set_error(const char *, ...) {}
int check_launch(const char *) { return 0; }
}  // namespace cn
using namespace cn;

__global__ void __launch_bounds__(576, 1)
ubench(int mode, int n_warps, int reps, int n_mma_bg, long long *out, float *sink) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[4];
  __shared__ uint32_t tmem_slot;
  __shared__ uint32_t stop_flag;
  if (threadIdx.x == 0) stop_flag = 0;
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t s_base = smem_u32(smem);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bars[i]), 1);
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
  fence_proxy_async_smem();
  if (warp == 0) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  constexpr uint32_t idesc = umma_idesc(128, 128, 0, 0);
  constexpr uint32_t idesc256 = umma_idesc(128, 256, 0, 0);
  long long t0 = 0, t1 = 0;

  if (mode == 0 || mode == 4) {
    if (threadIdx.x == 0) {
      t0 = clock64();
      for (int r = 0; r < reps; ++r) {
        for (int j = 0; j < 8; ++j)
          umma_f16(tmem, umma_desc(s_base + (j & 3) * 32 + (j >> 2) * 16384, 16, 1024),
                   umma_desc(s_base + 32768 + (j & 3) * 32 + (j >> 2) * (mode == 4 ? 32768 : 16384), 16, 1024),
                   mode == 4 ? idesc256 : idesc, 1u);
      }
      umma_commit(smem_u32(&bars[0]));
      mbar_wait(smem_u32(&bars[0]), 0);
      t1 = clock64();
      out[0] = (t1 - t0);
    }
  } else {
    // background MMA issuer: the last warp
    const bool bg = (n_mma_bg > 0) && warp == (uint32_t)n_warps;
    if (bg && lane == 0) {
      const long long tb = clock64();
      long long n = 0;
      uint32_t ph = 0;
      while (*reinterpret_cast<volatile uint32_t *>(&stop_flag) == 0) {
        for (int r = 0; r < n_mma_bg; ++r)
          for (int j = 0; j < 8; ++j)
            umma_f16(tmem + 384, umma_desc(s_base + 65536 + (j & 3) * 32 + (j >> 2) * 16384, 16, 1024),
                     umma_desc(s_base + 98304 + (j & 3) * 32 + (j >> 2) * 16384, 16, 1024), idesc, 1u);
        umma_commit(smem_u32(&bars[1]));
        mbar_wait(smem_u32(&bars[1]), ph);
        ph ^= 1;
        n += n_mma_bg * 8;
      }
      out[1] = clock64() - tb;
      out[2] = n;
    }
    if (warp < (uint32_t)n_warps) {
      // warps are grouped per 128x128 tile: n_split warps share a lane quadrant and split the columns
      const int n_split = n_warps >= 8 ? 2 : 1;            // column halves
      const int tile_warps = 4 * n_split;
      const uint32_t tile = warp / tile_warps;              // concurrent tiles
      const uint32_t quad = warp & 3u;
      const uint32_t half = (warp % tile_warps) >> 2;
      const uint32_t cols = 128 / n_split;
      const uint32_t row = quad * 32 + lane;
      const uint32_t taddr = tmem + ((quad * 32u) << 16) + (tile & 1) * 128 + half * cols;
      uint8_t *act = smem + 0 + (tile & 1) * 32768;
      const float *bias = reinterpret_cast<const float *>(smem + 140 * 1024);
      float acc = 0.f;
      __syncwarp();
      named_bar_sync(1, n_warps * 32);
      t0 = clock64();
      for (int r = 0; r < reps; ++r) {
        for (uint32_t c = 0; c < cols; c += 64) {
          uint32_t v[64];
          tmem_ld32(taddr + c, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
          tmem_ld32(taddr + c + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
          tmem_wait_ld();
          if (mode == 1) {
#pragma unroll
            for (int i = 0; i < 64; i += 16) acc += __uint_as_float(v[i]);
          } else if (mode == 5) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              uint32_t o[4];
#pragma unroll
              for (int i = 0; i < 4; ++i)
                asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(o[i]) : "f"(__uint_as_float(v[q * 8 + 2 * i + 1])), "f"(__uint_as_float(v[q * 8 + 2 * i])));
              *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, (half * cols + c) / 8 + q)) = make_uint4(o[0], o[1], o[2], o[3]);
            }
          } else if (mode == 6) {
            // bias add folded as one FFMA-free add with a per-thread register bias (transposed layout: lane = feature)
            const float bl = bias[row];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              uint32_t o[4];
#pragma unroll
              for (int i = 0; i < 4; ++i)
                asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(o[i]) : "f"(__uint_as_float(v[q * 8 + 2 * i + 1]) + bl), "f"(__uint_as_float(v[q * 8 + 2 * i]) + bl));
              *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, (half * cols + c) / 8 + q)) = make_uint4(o[0], o[1], o[2], o[3]);
            }
          } else {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 b0 = *reinterpret_cast<const float4 *>(bias + half * cols + c + q * 8);
              const float4 b1 = *reinterpret_cast<const float4 *>(bias + half * cols + c + q * 8 + 4);
              float f[8];
              f[0] = fmaxf(__uint_as_float(v[q * 8 + 0]) + b0.x, 0.f); f[1] = fmaxf(__uint_as_float(v[q * 8 + 1]) + b0.y, 0.f);
              f[2] = fmaxf(__uint_as_float(v[q * 8 + 2]) + b0.z, 0.f); f[3] = fmaxf(__uint_as_float(v[q * 8 + 3]) + b0.w, 0.f);
              f[4] = fmaxf(__uint_as_float(v[q * 8 + 4]) + b1.x, 0.f); f[5] = fmaxf(__uint_as_float(v[q * 8 + 5]) + b1.y, 0.f);
              f[6] = fmaxf(__uint_as_float(v[q * 8 + 6]) + b1.z, 0.f); f[7] = fmaxf(__uint_as_float(v[q * 8 + 7]) + b1.w, 0.f);
              *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, (half * cols + c) / 8 + q)) =
                  make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
            }
          }
        }
      }
      __syncwarp();
      named_bar_sync(1, n_warps * 32);
      t1 = clock64();
      if (threadIdx.x == 0) { out[0] = t1 - t0; *reinterpret_cast<volatile uint32_t *>(&stop_flag) = 1; }
      if (acc == 123.456f) sink[threadIdx.x] = acc;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

int main() {
  long long *out;
  float *sink;
  cudaMallocManaged(&out, 64);
  cudaMalloc(&sink, 4096);
  const size_t smem = 200 * 1024;
  cudaFuncSetAttribute(ubench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  auto run = [&](int mode, int nw, int reps, int bg) {
    out[0] = out[1] = 0; out[2] = 1;
    int threads = (nw + 1) * 32;
    if (threads < 128) threads = 128;
    ubench<<<1, threads, smem>>>(mode, nw, reps, bg, out, sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("mode %d: %s\n", mode, cudaGetErrorString(e)); exit(1); }
    return out[0];
  };
  for (int w = 0; w < 2; ++w) {  // second pass = warm
    long long c = run(0, 0, 64, 0);
    printf("mma 128x128x16: %.1f cyc/MMA (%d MMAs)\n", (double)c / (64 * 8), 64 * 8);
    c = run(4, 0, 64, 0);
    printf("mma 128x256x16: %.1f cyc/MMA\n", (double)c / (64 * 8));
    for (int nw : {4, 8, 16}) {
      const int tiles = nw >= 8 ? nw / 8 : 1;
      c = run(1, nw, 256, 0);
      printf("tmem_ld only, %2d warps: %.1f cyc per 128x128 fp32 tile (%d concurrent)\n", nw, (double)c / (256.0 * tiles), tiles);
      c = run(2, nw, 256, 0);
      printf("full epilogue, %2d warps: %.1f cyc per tile\n", nw, (double)c / (256.0 * tiles));
      c = run(2, nw, 256, 4);
      printf("full epilogue + bg MMA, %2d warps: %.1f cyc per tile; bg %.1f cyc/MMA\n", nw,
             (double)c / (256.0 * tiles), (double)out[1] / (double)out[2]);
      c = run(5, nw, 256, 0);
      printf("cvt.relu epilogue (no bias), %2d warps: %.1f cyc per tile\n", nw, (double)c / (256.0 * tiles));
      c = run(5, nw, 256, 4);
      printf("cvt.relu epilogue + bg MMA, %2d warps: %.1f cyc per tile; bg %.1f cyc/MMA\n", nw,
             (double)c / (256.0 * tiles), (double)out[1] / (double)out[2]);
      c = run(6, nw, 256, 0);
      printf("lane-bias cvt.relu epilogue, %2d warps: %.1f cyc per tile\n", nw, (double)c / (256.0 * tiles));
      c = run(1, nw, 256, 4);
      printf("tmem_ld only + bg MMA, %2d warps: %.1f cyc per tile; bg %.1f cyc/MMA\n", nw,
             (double)c / (256.0 * tiles), (double)out[1] / (double)out[2]);
    }
  }
  return 0;
}
