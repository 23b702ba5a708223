// How fast can ONE thread issue tcgen05.mma 128x128x16 with run-time operand addresses (ring slots), and what does
// the instruction sequence around each MMA cost?  Three issue styles, each with 0 or 16 busy ALU warps beside it:
//   style 0: lane 0 of a warp, descriptor rebuilt from the address per MMA  (umma_desc(addr + j * 32))
//   style 1: lane 0 of a warp, descriptor built once per slab, K step = +2 on the encoded low word
//   style 2: the whole warp runs the loop (warp-uniform control flow), the MMA under elect.sync, style-1 descriptors
// Reported: cycles per MMA until the last one is ISSUED, and until the last one has COMPLETED (commit -> mbarrier).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_mma_issue tools/ubench_mma_issue.cu
#include <cstdio>
#include <cstdlib>
#include "../thesis-clip-nerf_b200/csrc/common.cuh"

namespace cn {
void set_error(const char *, ...) {}
int check_launch(const char *) { return 0; }
}  // namespace cn
using namespace cn;

__global__ void __launch_bounds__(608, 1)
ubench(int style, int n_busy, int reps, int mul, long long *out, float *sink) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ uint32_t stop_flag;
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t s_base = smem_u32(smem);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    stop_flag = 0;
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 192 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
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
  constexpr int kBatch = 48;   // MMAs per batch = 12 (slab A, slab B) pairs x 4 K steps

  if (warp == 1) {
    if (style < 2) {
      if (lane == 0) {
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
          for (int i = 0; i < kBatch / 4; ++i) {
            const uint32_t a_addr = s_base + (((r * 12 + i) * mul + 3) & 3) * CN_SLAB_BYTES;          // run-time ring slot
            const uint32_t b_addr = s_base + (4 + (((r * 12 + i + 7) * mul) & 3)) * CN_SLAB_BYTES;
            if (style == 0) {
#pragma unroll
              for (int j = 0; j < 4; ++j)
                umma_f16(tmem + (i & 1) * 128, umma_desc(a_addr + j * kKmajKstep, kKmajLbo, kSboBytes),
                         umma_desc(b_addr + j * kKmajKstep, kKmajLbo, kSboBytes), idesc, 1u);
            } else {
              const uint64_t da = umma_desc(a_addr, kKmajLbo, kSboBytes), db = umma_desc(b_addr, kKmajLbo, kSboBytes);
#pragma unroll
              for (int j = 0; j < 4; ++j) umma_f16(tmem + (i & 1) * 128, da + 2 * j, db + 2 * j, idesc, 1u);
            }
          }
        }
        const long long t1 = clock64();
        umma_commit(smem_u32(&bar));
        mbar_wait(smem_u32(&bar), 0);
        const long long t2 = clock64();
        out[0] = t1 - t0;
        out[1] = t2 - t0;
        *reinterpret_cast<volatile uint32_t *>(&stop_flag) = 1;
      }
    } else {
      const long long t0 = clock64();
      for (int r = 0; r < reps; ++r) {
        for (int i = 0; i < kBatch / 4; ++i) {
          const uint32_t a_addr = s_base + (((r * 12 + i) * mul + 3) & 3) * CN_SLAB_BYTES;          // run-time ring slot
          const uint32_t b_addr = s_base + (4 + (((r * 12 + i + 7) * mul) & 3)) * CN_SLAB_BYTES;
          const uint64_t da = umma_desc(a_addr, kKmajLbo, kSboBytes), db = umma_desc(b_addr, kKmajLbo, kSboBytes);
          if (elect_one()) {
#pragma unroll
            for (int j = 0; j < 4; ++j) umma_f16(tmem + (i & 1) * 128, da + 2 * j, db + 2 * j, idesc, 1u);
          }
          __syncwarp();
        }
      }
      const long long t1 = clock64();
      if (elect_one()) umma_commit(smem_u32(&bar));
      __syncwarp();
      mbar_wait(smem_u32(&bar), 0);
      const long long t2 = clock64();
      if (lane == 0) {
        out[0] = t1 - t0;
        out[1] = t2 - t0;
        *reinterpret_cast<volatile uint32_t *>(&stop_flag) = 1;
      }
    }
  } else if (warp >= 2 && warp < 2 + (uint32_t)n_busy) {
    // busy ALU warps (an epilogue's worth of FMAs), 4 per scheduler when n_busy = 16
    float a = (float)threadIdx.x, b = 1.0001f;
    while (*reinterpret_cast<volatile uint32_t *>(&stop_flag) == 0) {
#pragma unroll
      for (int i = 0; i < 64; ++i) a = fmaf(a, b, 0.5f);
    }
    if (a == 123.456f) sink[threadIdx.x] = a;
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
  const int reps = 32;
  for (int pass = 0; pass < 2; ++pass)
    for (int busy : {0, 16})
      for (int style = 0; style < 3; ++style) {
        out[0] = out[1] = 0;
        ubench<<<1, 608, smem>>>(style, busy, reps, 5, out, sink);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("style %d: %s\n", style, cudaGetErrorString(e)); return 1; }
        if (pass)
          printf("style %d, %2d busy warps: %.1f cycles per MMA issued, %.1f per MMA completed (%d MMAs)\n", style, busy,
                 (double)out[0] / (reps * 48), (double)out[1] / (reps * 48), reps * 48);
      }
  return 0;
}
