// Shared device helpers for the sm_100a kernels: PTX wrappers (mbarrier, bulk-copy TMA,
// tcgen05 / TMEM), the 128-byte-swizzled bf16 tile addressing and host-side error plumbing.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <stdint.h>

#include "../../include/clipnerf_b200.h"

namespace cn {

// ---------------------------------------------------------------- host error plumbing
void set_error(const char *fmt, ...);
int check_launch(const char *what);
int check_scene(const CnScene *s, const char *who);

#define CN_REQUIRE(cond, ...)        \
  do {                               \
    if (!(cond)) {                   \
      ::cn::set_error(__VA_ARGS__);  \
      return -1;                     \
    }                                \
  } while (0)

// SM count of the current device (cudaDevAttrMultiProcessorCount, queried once per device): every persistent
// grid is sized from it, so a MIG slice or another SKU gets its own grid instead of a silent 148
int num_sms();

// ---------------------------------------------------------------- tile addressing
// byte offset of element (row, col) inside a bf16 tile image made of [128 x 64] slabs
__host__ __device__ __forceinline__ uint32_t tile_off(uint32_t row, uint32_t col) {
  return (col >> 6) * CN_SLAB_BYTES + row * 128u + ((((col >> 3) & 7u) ^ (row & 7u)) << 4) + (col & 7u) * 2u;
}
// byte offset of 16-byte chunk `chunk` (= col / 8) of `row`
__host__ __device__ __forceinline__ uint32_t tile_chunk_off(uint32_t row, uint32_t chunk) {
  return (chunk >> 3) * CN_SLAB_BYTES + row * 128u + (((chunk & 7u) ^ (row & 7u)) << 4);
}

// internal column order of the MLP input X  ->  reference column (layers.py:362), -1 = zero pad
__host__ __device__ __forceinline__ int x_col_to_ref(int j) {
  if (j < 256) return 123 + j;          // features
  if (j < 316) return j - 256;          // posenc(position)
  if (j < 376) return 60 + (j - 316);   // posenc(direction)
  if (j < 379) return 120 + (j - 376);  // 2*img-1
  return -1;
}

#ifdef __CUDACC__
// ---------------------------------------------------------------- optional per-role timeline (make TRACE=1)
// CTA 0 stamps clock64() into slot (iteration * 32 + point); every translation unit has its own buffer and
// exports a reader (CN_TRACE_READER), used by tools/trace_*.py to see where a warp-specialised pipeline waits.
#ifdef CN_TRACE
static __device__ long long cn_trace_buf[32 * 32];
#define CN_TRACE_AT(it, k)                                                              \
  do {                                                                                  \
    if (blockIdx.x == 0 && (it) < 32) cn_trace_buf[(it) * 32 + (k)] = clock64();        \
  } while (0)
#define CN_TRACE_VAL(it, k, v)                                                          \
  do {                                                                                  \
    if (blockIdx.x == 0 && (it) < 32) cn_trace_buf[(it) * 32 + (k)] = (long long)(v);   \
  } while (0)
#ifdef CN_TRACE_FINE   // per-epilogue accumulators: many clock reads in one warp, the traced tile runs ~30 % slower
#define CN_TRACE_CLOCK_IF(cond) ((cond) ? clock64() : 0ll)
#else
#define CN_TRACE_CLOCK_IF(cond) 0ll
#endif
#define CN_TRACE_READER(name)                                                           \
  extern "C" int name(long long *dst, int n) {                                          \
    return (int)cudaMemcpyFromSymbol(dst, cn_trace_buf, sizeof(long long) * n);         \
  }
#else
#define CN_TRACE_AT(it, k) do { } while (0)
#define CN_TRACE_VAL(it, k, v) do { } while (0)
#define CN_TRACE_CLOCK_IF(cond) 0ll
#define CN_TRACE_READER(name)
#endif

// ---------------------------------------------------------------- small utilities
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t *>(&v);
}
// packed bf16x2 helpers (low half = first argument / lower column)
__device__ __forceinline__ uint32_t cvt_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t cvt_relu_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// per half: 1.0 where a > 0 else 0.0
__device__ __forceinline__ uint32_t gt0_bf16x2(uint32_t a) {
  uint32_t r;
  asm("set.gt.bf16x2.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(0u));
  return r;
}
__device__ __forceinline__ uint32_t mul_bf16x2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t fma_bf16x2(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug becomes a trap (launch failure) instead of a hung GPU.
#ifndef CN_WATCHDOG_NS
#define CN_WATCHDOG_NS 4000000000ull
#endif
// The %globaltimer read in every failed poll doubles as the back-off of the wait.  Measured alternatives (round 2): reading
// the clock only every 4096th poll (tight polling) made the MLP kernels 5 ... 13 % SLOWER -- sixteen polling warps take
// issue slots and shared-memory bandwidth from the MMA thread and the working epilogue warps; try_wait with a 2 us
// suspend-time hint was no faster than this loop either; and 48 quick polls before this loop for the ONE MMA-issuing
// thread alone (its first poll often fails by a few cycles) made cn_mlp_block_bwd 2 % slower.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const uint64_t t0 = globaltimer_ns();
  while (!mbar_try_wait(bar, parity)) {
    if (globaltimer_ns() - t0 > CN_WATCHDOG_NS) {
      printf("clipnerf_b200: mbarrier watchdog block %d thread %d bar 0x%x parity %u\n", (int)blockIdx.x,
             (int)threadIdx.x, bar, parity);
      __trap();
    }
  }
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// generic-proxy writes to shared memory -> visible to the async proxy (TMA, tcgen05.mma)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---------------------------------------------------------------- bulk copies (TMA, 1-D)
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   dst_smem),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// global -> L2 only: starts the DRAM fetch of a range whose shared-memory slot is not free yet (bytes % 16 == 0)
__device__ __forceinline__ void bulk_prefetch_l2(const void *src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes)
               : "memory");
}
// shared -> global element-wise fp32 add performed by the TMA unit / L2 (no SM issue slots)
__device__ __forceinline__ void bulk_reduce_add_f32(void *dst, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst),
               "r"(src_smem), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ---------------------------------------------------------------- tcgen05 / TMEM
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start>>4 [0,14), LBO>>4 [16,30),
// SBO>>4 [32,46), version=1 [46,48), layout type [61,64) (2 = SWIZZLE_128B)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor), kind::f16, bf16 x bf16 -> f32
__host__ __device__ constexpr uint32_t umma_idesc(uint32_t M, uint32_t N, uint32_t a_mn, uint32_t b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
// D[tmem] (+)= A[smem] . B[smem]; issued by ONE thread
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once every previously issued tcgen05.mma of this thread has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// operand kinds used by the MLP kernels (all bf16, 128-byte swizzle, slab images):
//  K-major  (row = M/N index, 64 contiguous K per slab): SBO = 1024 (8-row group), K step 16 = +32 B
//  MN-major (row = K index, 64 contiguous M/N per slab): LBO = slab stride, SBO = 1024, K step 16 = +2048 B
constexpr uint32_t kSboBytes = 1024;
constexpr uint32_t kKmajLbo = 16;
constexpr uint32_t kKmajKstep = 32;
constexpr uint64_t kKmajKstepEnc = kKmajKstep >> 4;   // the same step on an encoded descriptor (start address field, no carry: smem < 256 KiB)
constexpr uint32_t kMNmajLbo = CN_SLAB_BYTES;
constexpr uint32_t kMNmajKstep = 2048;

// 32 lanes x 32 consecutive 32-bit columns: thread <-> TMEM lane (row), r[j] <-> column j
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}

// ---------------------------------------------------------------- global memory helpers
__device__ __forceinline__ float4 ldg_nc_f4(const float *p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ uint4 ldg_nc_u4(const void *p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void red_add_f4(float *p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
#endif  // __CUDACC__

}  // namespace cn
