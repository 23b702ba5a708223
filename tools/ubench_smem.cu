// Bank-conflict probe of the epilogue's shared-memory access pattern (uint4 chunks of a 128-byte-swizzled tile image):
// ncu --metrics l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum,
//               l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum ./ubench_smem
// variant 0: thread = (row = 32 quad + lane, column quarter), chunk c ^ (row & 7) -- the block backward's E0/E1/E2
// variant 1: the same without the swizzle (every lane of a quarter-warp on the same 16-byte column: 8-way conflicts)
// variant 2: the forward's mapping (8 chunks of one column half per thread)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t off(uint32_t row, uint32_t chunk, int swz) {
  return (chunk >> 3) * 16384u + row * 128u + ((((chunk & 7u) ^ (swz ? (row & 7u) : 0u))) << 4);
}
template <int V> __global__ void __launch_bounds__(512) probe(uint4 *out, int iters) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t quad = warp & 3u, cq = warp >> 2, row = quad * 32u + lane;
  uint4 acc = make_uint4(threadIdx.x, 1, 2, 3);
  for (int it = 0; it < iters; ++it) {
    if (V == 2) {
      const uint32_t half = cq & 1u;
#pragma unroll
      for (int q = 0; q < 8; ++q) *reinterpret_cast<uint4 *>(smem + off(row, half * 8 + q, 1)) = acc;
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 *p = reinterpret_cast<uint4 *>(smem + off(row, cq * 4 + q, V == 0));
        const uint4 v = *p;
        acc.x += v.x; acc.y ^= v.y;
        *p = acc;
      }
    }
    __syncthreads();
  }
  out[threadIdx.x] = acc;
}
int main() {
  uint4 *d;
  cudaMalloc(&d, 512 * 16);
  cudaFuncSetAttribute(probe<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(probe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  probe<0><<<1, 512, 65536>>>(d, 100);
  probe<1><<<1, 512, 65536>>>(d, 100);
  probe<2><<<1, 512, 65536>>>(d, 100);
  printf("rc %d\n", (int)cudaDeviceSynchronize());
  return 0;
}
