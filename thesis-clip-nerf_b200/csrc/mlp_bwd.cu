// Backward of one pre-activation ResNet block (layers.py:262-298) for every 128-row tile, on
// tcgen05 / TMEM:      y = h + relu(relu(h) Wa + ba) Wb + bb
//   saved by the forward: A1 = relu(h) only (bf16 tile image).  The hidden activation is RECOMPUTED
//   here -- one extra GEMM per tile buys a third of the kernel's HBM traffic, which is what bounds it:
//   T  : t    = A1 Wa             (acc0; B = the dgrad image of Wa read MN-major: bit-identical to the forward)
//   E0 : A2   = relu(t + ba)      -> bf16 operand tile in smem
//   G2 : g_a2 = g_y Wb^T          (acc1, K-major operands)          issued while E0 runs
//   WB : dWb += A2^T g_y          (MN-major views of the same smem tiles)
//   E1 : g_t  = g_a2 * [A2 > 0]   -> overwrites A2 in place (same thread, same bytes)
//   G1 : g_a1 = g_t Wa^T          (acc1)
//   WA : dWa += A1^T g_t
//   E2 : g_h  = g_y + g_a1 * [A1 > 0]  -> overwrites g_y in place, TMA store to global
//   the next tile's T is issued behind WB so that its E0 never waits for the tensor pipe.
// Cost model (tools/ubench_tc.cu, tools/trace_block_bwd.py, B200): MMA 128x128x16 = 64 cycles (512 per GEMM), and an
// SS-mode 128x128 MMA reads its 8 KiB of operands at 128 B/clk -- ALL of the SM's shared-memory bandwidth.  The five
// GEMMs of a tile already move 320 KiB through shared memory, so the kernel is bound by SHARED-MEMORY TRAFFIC before
// it is bound by the tensor pipe or by its 96 KiB of HBM traffic per tile.  Hence:
//   * every epilogue is 3 packed instructions per element PAIR (cvt.bf16x2 / set.bf16x2 / mul|fma.bf16x2), 16 warps;
//   * the ReLU mask of A2 stays in registers between E0 and E1 (no re-read of the tile);
//   * the bias gradients (column sums of g_t and g_y) are taken from the registers the epilogue already holds, by a
//     packed bf16x2 butterfly over the 32 rows of a warp (15 shuffles), accumulated in fp32 across tiles -- no extra
//     pass over the tiles in shared memory;
//   * g_h replaces g_y in its buffer and leaves by ONE bulk store (a store straight from registers costs 32 partial
//     lines per instruction: measured 2000 cycles per tile in the LSU).
// Shared memory: Wa, Wb (dgrad images, resident) + 2 x A1 + 2 x g_y (double buffered TMA) + 1 x A2|g_t.
// TMEM: acc0, acc1, and the two weight gradients, resident across all tiles of the CTA and flushed once
// with vector red.global.add.
//   warp 0: TMA producer, warp 1: MMA issuer, warps 2-17: epilogue (4 TMEM lane quadrants x 4 column quarters).
#include "common.cuh"
#include "mlp_layout.cuh"

namespace cn {

constexpr int kBwdEpiWarps = 16;
constexpr int kBwdEpiThreads = kBwdEpiWarps * 32;
constexpr int kBwdThreads = 64 + kBwdEpiThreads;
constexpr uint32_t kBwSmemWb = 0;                       // Wb, dgrad image
constexpr uint32_t kBwSmemWa = 1 * CN_ACT_TILE_BYTES;   // Wa, dgrad image (also the recompute's MN-major B)
constexpr uint32_t kBwSmemA1 = 2 * CN_ACT_TILE_BYTES;   // 2 buffers
constexpr uint32_t kBwSmemGy = 4 * CN_ACT_TILE_BYTES;   // 2 buffers (g_y in, g_h out)
constexpr uint32_t kBwSmemA2 = 6 * CN_ACT_TILE_BYTES;   // A2, then g_t
constexpr uint32_t kBwdSmemBytes = 7 * CN_ACT_TILE_BYTES + 1024;

struct BwdBars {
  uint64_t w_full;
  uint64_t full_a1[2], full_gy[2];     // TMA landed
  // epilogue / bias warps arrive ONCE PER WARP (lane 0 behind __syncwarp): 16 resp. 2 arrivals
  uint64_t free_a1[2];                 // E2 warps + commit behind WA
  uint64_t gy_staged[2];               // E2 warps: g_h has replaced g_y in its buffer (G2 / WB are behind G1: done)
  uint64_t t_ready, g2wb_done, g1_ready;  // commits
  uint64_t a2_ready, gt_ready;         // epilogue warps
  uint64_t a2buf_free;                 // commit behind WA
  uint64_t acc1_free;                  // E2 warps
  uint64_t all_done;
};

// Column sums over the 32 rows (lanes) of a warp of 16 packed bf16x2 words per lane (32 columns): a butterfly that
// halves what a lane holds at every step.  The first two levels (sums of 2 and 4 values) stay packed bf16x2
// (8 + 4 exchanges), the last three run in fp32 (4 + 2 + 1 exchanges) -- the rounding of a packed add grows with the
// partial sum, so the late levels are the ones worth 32 bits.  Returns, in lane l, the sum of column l.
__device__ __forceinline__ float colsum_bf16x2(uint32_t (&w)[16], uint32_t lane) {
#pragma unroll
  for (int lvl = 0; lvl < 2; ++lvl) {
    const int half = 8 >> lvl;                 // words kept after this level
    const uint32_t bit = 16u >> lvl;           // partner = lane ^ bit; lanes with the bit set keep the upper half
    const bool up = (lane & bit) != 0;
#pragma unroll
    for (int j = 0; j < half; ++j) {
      const uint32_t send = up ? w[j] : w[j + half];
      const uint32_t keep = up ? w[j + half] : w[j];
      const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, bit);
      asm("add.rn.bf16x2 %0, %1, %2;" : "=r"(w[j]) : "r"(keep), "r"(recv));
    }
  }
  float f[8];  // words 4 (lane >> 3) + {0..3}  ->  columns 8 (lane >> 3) + {0..7}
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    f[2 * j] = bf16_lo(w[j]);
    f[2 * j + 1] = bf16_hi(w[j]);
  }
#pragma unroll
  for (int lvl = 0; lvl < 3; ++lvl) {
    const int half = 4 >> lvl;
    const uint32_t bit = 4u >> lvl;
    const bool up = (lane & bit) != 0;
#pragma unroll
    for (int j = 0; j < half; ++j) {
      const float send = up ? f[j] : f[j + half];
      const float keep = up ? f[j + half] : f[j];
      f[j] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
    }
  }
  return f[0];
}

__global__ void __launch_bounds__(kBwdThreads, 1)
mlp_block_bwd_kernel(const uint8_t *__restrict__ w_a_img, const uint8_t *__restrict__ w_b_img,
                     const float *__restrict__ bias_a, const uint8_t *__restrict__ acts, int64_t act_stride_bytes,
                     int64_t n_tiles, const uint8_t *__restrict__ gy_tiles, int64_t gy_views, int64_t gy_tpv,
                     uint8_t *__restrict__ gh_tiles, float scale, float *__restrict__ g_wa, float *__restrict__ g_ba,
                     float *__restrict__ g_wb, float *__restrict__ g_bb) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) BwdBars bars;
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float s_ba[128];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t s_base = smem_u32(smem);
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();

  if (threadIdx.x < 128) s_ba[threadIdx.x] = bias_a[threadIdx.x];
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bars.w_full), 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(smem_u32(&bars.full_a1[b]), 1);
      mbar_init(smem_u32(&bars.full_gy[b]), 1);
      mbar_init(smem_u32(&bars.free_a1[b]), kBwdEpiWarps + 1);
      mbar_init(smem_u32(&bars.gy_staged[b]), kBwdEpiWarps);
    }
    mbar_init(smem_u32(&bars.t_ready), 1);
    mbar_init(smem_u32(&bars.g2wb_done), 1);
    mbar_init(smem_u32(&bars.g1_ready), 1);
    mbar_init(smem_u32(&bars.a2_ready), kBwdEpiWarps);
    mbar_init(smem_u32(&bars.gt_ready), kBwdEpiWarps);
    mbar_init(smem_u32(&bars.a2buf_free), 1);
    mbar_init(smem_u32(&bars.acc1_free), kBwdEpiWarps);
    mbar_init(smem_u32(&bars.all_done), 1);
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
  const uint32_t t_acc0 = tmem, t_acc1 = tmem + 128, t_dwa = tmem + 256, t_dwb = tmem + 384;
  const int64_t n_mine = n_tiles > blockIdx.x ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      mbar_expect_tx(smem_u32(&bars.w_full), 2 * CN_ACT_TILE_BYTES);
      bulk_g2s(s_base + kBwSmemWb, w_b_img, CN_ACT_TILE_BYTES, smem_u32(&bars.w_full));
      bulk_g2s(s_base + kBwSmemWa, w_a_img, CN_ACT_TILE_BYTES, smem_u32(&bars.w_full));
      auto load_a1 = [&](int64_t it) {
        const int64_t tile = blockIdx.x + it * gridDim.x;
        const uint32_t b = (uint32_t)it & 1u;
        mbar_expect_tx(smem_u32(&bars.full_a1[b]), CN_ACT_TILE_BYTES);
        bulk_g2s(s_base + kBwSmemA1 + b * CN_ACT_TILE_BYTES, acts + tile * act_stride_bytes, CN_ACT_TILE_BYTES,
                 smem_u32(&bars.full_a1[b]));
      };
      auto load_gy = [&](int64_t it) {
        const int64_t tile = blockIdx.x + it * gridDim.x;
        const uint32_t b = (uint32_t)it & 1u;
        mbar_expect_tx(smem_u32(&bars.full_gy[b]), CN_ACT_TILE_BYTES);
        // mean over views folded in: every view tile of a (scene, row tile) reads the same fusion-stage g tile
        const int64_t src = gy_views > 1 ? (tile / (gy_views * gy_tpv)) * gy_tpv + tile % gy_tpv : tile;
        bulk_g2s(s_base + kBwSmemGy + b * CN_ACT_TILE_BYTES, gy_tiles + src * (int64_t)CN_ACT_TILE_BYTES,
                 CN_ACT_TILE_BYTES, smem_u32(&bars.full_gy[b]));
      };
      for (int64_t it = 0; it < 2 && it < n_mine; ++it) {
        load_a1(it);
        load_gy(it);
      }
      for (int64_t it = 0; it < n_mine; ++it) {
        const int64_t tile = blockIdx.x + it * gridDim.x;
        const uint32_t b = (uint32_t)it & 1u, ph2 = ((uint32_t)it >> 1) & 1u;
        // E2 of this tile has replaced g_y by g_h in place: stream it out, then refill the buffer
        mbar_wait(smem_u32(&bars.gy_staged[b]), ph2);
        CN_TRACE_AT(it, 24);
        bulk_s2g(gh_tiles + tile * (int64_t)CN_ACT_TILE_BYTES, s_base + kBwSmemGy + b * CN_ACT_TILE_BYTES,
                 CN_ACT_TILE_BYTES);
        bulk_commit();
        if (it + 2 < n_mine) {
          bulk_wait_read0();
          load_gy(it + 2);
          mbar_wait(smem_u32(&bars.free_a1[b]), ph2);
          CN_TRACE_AT(it, 26);
          load_a1(it + 2);
        }
      }
      bulk_wait0();
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      constexpr uint32_t idesc_kk = umma_idesc(128, 128, 0, 0);  // dgrad: both operands K-major
      constexpr uint32_t idesc_kn = umma_idesc(128, 128, 0, 1);  // recompute: A K-major, B MN-major
      constexpr uint32_t idesc_mn = umma_idesc(128, 128, 1, 1);  // wgrad: both operands MN-major
      auto gemm_kk = [&](uint32_t d_tmem, uint32_t a_off, uint32_t b_off) {
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
          for (int j = 0; j < 4; ++j)
            umma_f16(d_tmem, umma_desc(s_base + a_off + s * CN_SLAB_BYTES + j * kKmajKstep, kKmajLbo, kSboBytes),
                     umma_desc(s_base + b_off + s * CN_SLAB_BYTES + j * kKmajKstep, kKmajLbo, kSboBytes), idesc_kk,
                     (s | j) ? 1u : 0u);
      };
      // t = A1 Wa with Wa stored [in][out] (K rows, N contiguous): K step 16 = 16 rows of the weight image
      auto gemm_t = [&](uint32_t a_off) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          umma_f16(t_acc0, umma_desc(s_base + a_off + (j >> 2) * CN_SLAB_BYTES + (j & 3) * kKmajKstep, kKmajLbo, kSboBytes),
                   umma_desc(s_base + kBwSmemWa + j * kMNmajKstep, kMNmajLbo, kSboBytes), idesc_kn, j ? 1u : 0u);
      };
      auto wgrad = [&](uint32_t d_tmem, uint32_t a_off, uint32_t b_off, bool first) {
#pragma unroll
        for (int j = 0; j < 8; ++j)  // 8 steps of 16 rows
          umma_f16(d_tmem, umma_desc(s_base + a_off + j * kMNmajKstep, kMNmajLbo, kSboBytes),
                   umma_desc(s_base + b_off + j * kMNmajKstep, kMNmajLbo, kSboBytes), idesc_mn,
                   (!first || j) ? 1u : 0u);
      };
      mbar_wait(smem_u32(&bars.w_full), 0);
      if (n_mine > 0) {
        mbar_wait(smem_u32(&bars.full_a1[0]), 0);
        tc_fence_after();
        gemm_t(kBwSmemA1);
        umma_commit(smem_u32(&bars.t_ready));
      }
      for (int64_t it = 0; it < n_mine; ++it) {
        const uint32_t b = (uint32_t)it & 1u, ph2 = ((uint32_t)it >> 1) & 1u, par = (uint32_t)it & 1u;
        const uint32_t a1 = kBwSmemA1 + b * CN_ACT_TILE_BYTES, gy = kBwSmemGy + b * CN_ACT_TILE_BYTES;
        const bool has_next = it + 1 < n_mine;
        const uint32_t nb = b ^ 1u, nph2 = ((uint32_t)(it + 1) >> 1) & 1u;
        mbar_wait(smem_u32(&bars.full_gy[b]), ph2);
        mbar_wait(smem_u32(&bars.acc1_free), par ^ 1u);          // E2 of the previous tile has drained acc1
        tc_fence_after();
        CN_TRACE_AT(it, 0);
        gemm_kk(t_acc1, gy, kBwSmemWb);                           // G2: g_a2 = g_y Wb^T
        CN_TRACE_AT(it, 1);
        mbar_wait(smem_u32(&bars.a2_ready), par);                 // (also: E0 has drained acc0)
        tc_fence_after();
        CN_TRACE_AT(it, 2);
        wgrad(t_dwb, kBwSmemA2, gy, it == 0);                     // WB: dWb += A2^T g_y
        umma_commit(smem_u32(&bars.g2wb_done));
        CN_TRACE_AT(it, 3);
        bool next_issued = false;
        if (has_next && mbar_try_wait(smem_u32(&bars.full_a1[nb]), nph2)) {
          tc_fence_after();
          gemm_t(kBwSmemA1 + nb * CN_ACT_TILE_BYTES);             // T of the next tile, behind WB
          umma_commit(smem_u32(&bars.t_ready));
          next_issued = true;
        }
        CN_TRACE_AT(it, 4);
        mbar_wait(smem_u32(&bars.gt_ready), par);
        tc_fence_after();
        CN_TRACE_AT(it, 5);
        gemm_kk(t_acc1, kBwSmemA2, kBwSmemWa);                    // G1: g_a1 = g_t Wa^T
        umma_commit(smem_u32(&bars.g1_ready));
        CN_TRACE_AT(it, 6);
        wgrad(t_dwa, a1, kBwSmemA2, it == 0);                     // WA: dWa += A1^T g_t
        umma_commit(smem_u32(&bars.free_a1[b]));
        umma_commit(smem_u32(&bars.a2buf_free));
        CN_TRACE_AT(it, 7);
        if (has_next && !next_issued) {
          mbar_wait(smem_u32(&bars.full_a1[nb]), nph2);
          tc_fence_after();
          gemm_t(kBwSmemA1 + nb * CN_ACT_TILE_BYTES);
          umma_commit(smem_u32(&bars.t_ready));
        }
      }
      umma_commit(smem_u32(&bars.all_done));
    }
  } else {
    // =========================== epilogue: 16 warps, (row, 32-column quarter) per thread ===========================
    const uint32_t e = warp - 2;
    const uint32_t cq = e >> 2;                      // column quarter [32 cq, 32 cq + 32)
    const uint32_t quad = warp & 3u;                 // TMEM lane quadrant this warp may access
    const uint32_t row = quad * 32u + lane;
    const uint32_t lane_off = (quad * 32u) << 16;
    uint32_t choff[4];                               // byte offsets of this thread's four 16-byte chunks in a tile
#pragma unroll
    for (int q = 0; q < 4; ++q) choff[q] = tile_chunk_off(row, cq * 4 + q);
    uint8_t *const a2buf = smem + kBwSmemA2;
    // bias gradients: lane l accumulates column 32 cq + l, summed over the warp's 32 rows of every tile
    float dba = 0.f, dbb = 0.f;
    for (int64_t it = 0; it < n_mine; ++it) {
      const uint32_t b = (uint32_t)it & 1u, par = (uint32_t)it & 1u;
      const uint8_t *const a1buf = smem + kBwSmemA1 + b * CN_ACT_TILE_BYTES;
      uint8_t *const gybuf = smem + kBwSmemGy + b * CN_ACT_TILE_BYTES;
      uint32_t a2w[16];                               // A2 of this thread: ReLU mask source of E1
      // ---- E0: A2 = relu(t + ba) -> bf16 operand tile
      mbar_wait(smem_u32(&bars.t_ready), par);
      mbar_wait(smem_u32(&bars.a2buf_free), par ^ 1u);  // WA of the previous tile is done with g_t
      tc_fence_after();
      if (threadIdx.x == 64) CN_TRACE_AT(it, 8);
      {
        uint32_t r[32];
        tmem_ld32(t_acc0 + lane_off + cq * 32, r);
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 b0 = *reinterpret_cast<const float4 *>(s_ba + cq * 32 + q * 8);
          const float4 b1 = *reinterpret_cast<const float4 *>(s_ba + cq * 32 + q * 8 + 4);
          a2w[q * 4 + 0] = cvt_relu_bf16x2(__uint_as_float(r[q * 8 + 0]) + b0.x, __uint_as_float(r[q * 8 + 1]) + b0.y);
          a2w[q * 4 + 1] = cvt_relu_bf16x2(__uint_as_float(r[q * 8 + 2]) + b0.z, __uint_as_float(r[q * 8 + 3]) + b0.w);
          a2w[q * 4 + 2] = cvt_relu_bf16x2(__uint_as_float(r[q * 8 + 4]) + b1.x, __uint_as_float(r[q * 8 + 5]) + b1.y);
          a2w[q * 4 + 3] = cvt_relu_bf16x2(__uint_as_float(r[q * 8 + 6]) + b1.z, __uint_as_float(r[q * 8 + 7]) + b1.w);
          *reinterpret_cast<uint4 *>(a2buf + choff[q]) =
              make_uint4(a2w[q * 4 + 0], a2w[q * 4 + 1], a2w[q * 4 + 2], a2w[q * 4 + 3]);
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars.a2_ready));
      if (threadIdx.x == 64) CN_TRACE_AT(it, 11);
      // ---- E1: g_t = g_a2 * [A2 > 0], over A2 (WB has finished reading it); dba += colsum(g_t)
      mbar_wait(smem_u32(&bars.g2wb_done), par);
      tc_fence_after();
      if (threadIdx.x == 64) CN_TRACE_AT(it, 12);
      {
        uint32_t r[32];
        tmem_ld32(t_acc1 + lane_off + cq * 32, r);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 16; ++j)
          a2w[j] = mul_bf16x2(cvt_bf16x2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1])), gt0_bf16x2(a2w[j]));
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<uint4 *>(a2buf + choff[q]) =
              make_uint4(a2w[q * 4 + 0], a2w[q * 4 + 1], a2w[q * 4 + 2], a2w[q * 4 + 3]);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars.gt_ready));
      if (threadIdx.x == 64) CN_TRACE_AT(it, 14);
      {
        dba += colsum_bf16x2(a2w, lane);               // off the critical path: G1 is running
      }
      // ---- E2: g_h = g_y + g_a1 * [A1 > 0], in place over g_y (TMA store by the producer warp); dbb += colsum(g_y)
      uint32_t gw[16];
      {
        uint4 g[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) g[q] = *reinterpret_cast<const uint4 *>(gybuf + choff[q]);  // (landed before G2)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          gw[q * 4 + 0] = g[q].x; gw[q * 4 + 1] = g[q].y; gw[q * 4 + 2] = g[q].z; gw[q * 4 + 3] = g[q].w;
        }
      }
      {
        uint32_t t16[16];                              // the butterfly consumes its input: work on a copy, while G1 runs
#pragma unroll
        for (int j = 0; j < 16; ++j) t16[j] = gw[j];
        dbb += colsum_bf16x2(t16, lane);
      }
      mbar_wait(smem_u32(&bars.g1_ready), par);
      tc_fence_after();
      if (threadIdx.x == 64) CN_TRACE_AT(it, 15);
      {
        uint32_t r[32];
        tmem_ld32(t_acc1 + lane_off + cq * 32, r);
        uint4 a[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) a[q] = *reinterpret_cast<const uint4 *>(a1buf + choff[q]);
        tmem_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars.acc1_free));
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 o;
          o.x = fma_bf16x2(cvt_bf16x2(__uint_as_float(r[q * 8 + 0]), __uint_as_float(r[q * 8 + 1])), gt0_bf16x2(a[q].x), gw[q * 4 + 0]);
          o.y = fma_bf16x2(cvt_bf16x2(__uint_as_float(r[q * 8 + 2]), __uint_as_float(r[q * 8 + 3])), gt0_bf16x2(a[q].y), gw[q * 4 + 1]);
          o.z = fma_bf16x2(cvt_bf16x2(__uint_as_float(r[q * 8 + 4]), __uint_as_float(r[q * 8 + 5])), gt0_bf16x2(a[q].z), gw[q * 4 + 2]);
          o.w = fma_bf16x2(cvt_bf16x2(__uint_as_float(r[q * 8 + 6]), __uint_as_float(r[q * 8 + 7])), gt0_bf16x2(a[q].w), gw[q * 4 + 3]);
          *reinterpret_cast<uint4 *>(gybuf + choff[q]) = o;
        }
      }
      fence_proxy_async_smem();
      __syncwarp();                                  // every lane has consumed its A1 chunks and written g_h
      if (lane == 0) {
        mbar_arrive(smem_u32(&bars.free_a1[b]));
        mbar_arrive(smem_u32(&bars.gy_staged[b]));
      }
      if (threadIdx.x == 64) CN_TRACE_AT(it, 17);
    }
    // ---- flush the weight gradients held in TMEM (each thread: its input unit, 32 output units)
    mbar_wait(smem_u32(&bars.all_done), 0);
    tc_fence_after();
    if (n_mine > 0) {
#pragma unroll 1
      for (int mtx = 0; mtx < 2; ++mtx) {
        float *gdst = (mtx == 0 ? g_wa : g_wb) + row * 128 + cq * 32;  // accumulator row = input unit
        uint32_t r[32];
        tmem_ld32((mtx == 0 ? t_dwa : t_dwb) + lane_off + cq * 32, r);
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 8; ++q)
          red_add_f4(gdst + q * 4, __uint_as_float(r[q * 4 + 0]) * scale, __uint_as_float(r[q * 4 + 1]) * scale,
                     __uint_as_float(r[q * 4 + 2]) * scale, __uint_as_float(r[q * 4 + 3]) * scale);
      }
      // ---- bias gradients: the four row quadrants of a column meet in shared memory (the tiles are dead now),
      // one atomic per column and CTA
      float *red = reinterpret_cast<float *>(a2buf);   // [2 biases][4 quadrants][128 columns]
      red[(0 * 4 + quad) * 128 + cq * 32 + lane] = dba;
      red[(1 * 4 + quad) * 128 + cq * 32 + lane] = dbb;
      named_bar_sync(1, kBwdEpiThreads);
      const uint32_t et = threadIdx.x - 64;
      if (et < 256) {
        const uint32_t which = et >> 7, col = et & 127u;
        const float v = red[(which * 4 + 0) * 128 + col] + red[(which * 4 + 1) * 128 + col] +
                        red[(which * 4 + 2) * 128 + col] + red[(which * 4 + 3) * 128 + col];
        atomicAdd((which == 0 ? g_ba : g_bb) + col, v * scale);
      }
    }
  }

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
  CN_REQUIRE(packed && acts && gy_tiles && gh_tiles && grad, "cn_mlp_block_bwd: null pointer");
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
  cudaError_t e =
      cudaFuncSetAttribute(mlp_block_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBwdSmemBytes);
  if (e != cudaSuccess) {
    set_error("cn_mlp_block_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  const uint8_t *pk = (const uint8_t *)packed;
  const unsigned grid = (unsigned)(n_tiles < num_sms() ? n_tiles : num_sms());
  mlp_block_bwd_kernel<<<grid, kBwdThreads, kBwdSmemBytes, (cudaStream_t)stream>>>(
      pk + pack_bwd_block(blk, 0), pk + pack_bwd_block(blk, 1),
      reinterpret_cast<const float *>(pk + kPackTables) + (blk < 3 ? kTabFeature : kTabFusion) + (2 * (blk % 3) + 1) * 128,
      (const uint8_t *)acts + (int64_t)act_index * CN_ACT_TILE_BYTES, (int64_t)act_stride_tiles * CN_ACT_TILE_BYTES,
      n_tiles, (const uint8_t *)gy_tiles, (int64_t)gy_views, gy_tiles_per_view, (uint8_t *)gh_tiles, scale,
      grad + param_block_w(blk, 0),
      grad + param_block_b(blk, 0), grad + param_block_w(blk, 1), grad + param_block_b(blk, 1));
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
