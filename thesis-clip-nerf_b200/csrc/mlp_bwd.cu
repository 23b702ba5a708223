// Backward of one pre-activation ResNet block (layers.py:262-298) for every 128-row tile, on
// tcgen05 / TMEM:      y = h + relu(relu(h) Wa + ba) Wb + bb
//   saved by the forward: A1 = relu(h) only (bf16 tile image).  The hidden activation is RECOMPUTED
//   here -- one extra GEMM per tile buys a third of the kernel's HBM traffic, which is what bounds it:
//   A2   = relu(A1 Wa + ba)    (forward GEMM, same bf16 operands and K order as the forward: bit-identical)
//   g_a2 = g_y Wb^T            (dgrad 1, K-major operands)
//   dWb += A2^T g_y            (wgrad,  MN-major operands: the SAME smem tiles read transposed)
//   g_t  = g_a2 * [A2 > 0]     (epilogue 1)  -> bf16 tile in smem,  dba += colsum(g_t)   (bias warps)
//   g_a1 = g_t Wa^T            (dgrad 2)
//   dWa += A1^T g_t            (wgrad)
//   g_h  = g_y + g_a1 * [A1>0] (epilogue 2)  -> global,             dbb += colsum(g_y)   (bias warps)
// The three weight images stay resident in shared memory for the whole kernel and the two weight
// gradients stay resident in TMEM (2 x 128 columns) across all tiles of the CTA; they are
// flushed once with vector red.global.add at the end.
//   warp 0: TMA producer (g_y, A1 tiles), warp 1: MMA issuer, warps 2-9: epilogue (4 TMEM lane
//   quadrants x 2 column halves), warps 10-11: bias gradients (column sums read from the smem tiles).
#include "common.cuh"
#include "mlp_layout.cuh"

namespace cn {

constexpr int kBwdThreads = 384;  // warp 0 producer, 1 MMA, 2-9 epilogue (4 lane quadrants x 2 column halves), 10-11 bias sums
constexpr int kBwdEpiThreads = 256;
constexpr uint32_t kBwSmemWb = 0;                       // Wb, dgrad image
constexpr uint32_t kBwSmemWa = 1 * CN_ACT_TILE_BYTES;   // Wa, dgrad image
constexpr uint32_t kBwSmemWf = 2 * CN_ACT_TILE_BYTES;   // Wa, forward image (recompute of the hidden activation)
constexpr uint32_t kBwSmemGy = 3 * CN_ACT_TILE_BYTES;
constexpr uint32_t kBwSmemA1 = 4 * CN_ACT_TILE_BYTES;
constexpr uint32_t kBwSmemA2 = 5 * CN_ACT_TILE_BYTES;   // written by epilogue 0
constexpr uint32_t kBwSmemGt = 6 * CN_ACT_TILE_BYTES;   // written by epilogue 1
constexpr uint32_t kBwdSmemBytes = 7 * CN_ACT_TILE_BYTES + 1024;

struct BwdBars {
  uint64_t w_full;
  uint64_t full_gy, full_a1;
  uint64_t free_gy;                    // 256 epilogue + 64 bias-sum arrivals + 1 tcgen05.commit
  uint64_t free_a1;                    // 256 epilogue arrivals + 1 tcgen05.commit
  uint64_t t_ready;                    // commit: A1 Wa is in TMEM
  uint64_t a2_ready;                   // 256 epilogue arrivals: relu(t) is in shared memory
  uint64_t a2_free;                    // commit after wgrad-b
  uint64_t gt_ready;                   // 256 epilogue arrivals
  uint64_t gt_free;                    // commit after wgrad-a + 64 bias-sum arrivals
  uint64_t tmp1_ready, tmp2_ready;     // commits
  uint64_t all_done;                   // commit after the last tile
};

__global__ void __launch_bounds__(kBwdThreads, 1)
mlp_block_bwd_kernel(const uint8_t *__restrict__ w_a_img, const uint8_t *__restrict__ w_b_img,
                     const uint8_t *__restrict__ w_a_fwd_img, const float *__restrict__ bias_a,
                     const uint8_t *__restrict__ acts, int64_t act_stride_bytes, int64_t n_tiles,
                     const uint8_t *__restrict__ gy_tiles, uint8_t *__restrict__ gh_tiles, float *__restrict__ g_wa,
                     float *__restrict__ g_ba, float *__restrict__ g_wb, float *__restrict__ g_bb) {
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
    mbar_init(smem_u32(&bars.full_gy), 1);
    mbar_init(smem_u32(&bars.full_a1), 1);
    mbar_init(smem_u32(&bars.free_gy), kBwdEpiThreads + 64 + 1);
    mbar_init(smem_u32(&bars.free_a1), kBwdEpiThreads + 1);
    mbar_init(smem_u32(&bars.t_ready), 1);
    mbar_init(smem_u32(&bars.a2_ready), kBwdEpiThreads);
    mbar_init(smem_u32(&bars.a2_free), 1);
    mbar_init(smem_u32(&bars.gt_ready), kBwdEpiThreads);
    mbar_init(smem_u32(&bars.gt_free), 64 + 1);
    mbar_init(smem_u32(&bars.tmp1_ready), 1);
    mbar_init(smem_u32(&bars.tmp2_ready), 1);
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
  const uint32_t t_tmp1 = tmem, t_tmp2 = tmem + 128, t_dwa = tmem + 256, t_dwb = tmem + 384;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      mbar_expect_tx(smem_u32(&bars.w_full), 3 * CN_ACT_TILE_BYTES);
      bulk_g2s(s_base + kBwSmemWb, w_b_img, CN_ACT_TILE_BYTES, smem_u32(&bars.w_full));
      bulk_g2s(s_base + kBwSmemWa, w_a_img, CN_ACT_TILE_BYTES, smem_u32(&bars.w_full));
      bulk_g2s(s_base + kBwSmemWf, w_a_fwd_img, CN_ACT_TILE_BYTES, smem_u32(&bars.w_full));
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        const uint32_t par = (it & 1u) ^ 1u;
        mbar_wait(smem_u32(&bars.free_a1), par);
        mbar_expect_tx(smem_u32(&bars.full_a1), CN_ACT_TILE_BYTES);
        bulk_g2s(s_base + kBwSmemA1, acts + tile * act_stride_bytes, CN_ACT_TILE_BYTES, smem_u32(&bars.full_a1));
        mbar_wait(smem_u32(&bars.free_gy), par);
        mbar_expect_tx(smem_u32(&bars.full_gy), CN_ACT_TILE_BYTES);
        bulk_g2s(s_base + kBwSmemGy, gy_tiles + tile * (int64_t)CN_ACT_TILE_BYTES, CN_ACT_TILE_BYTES,
                 smem_u32(&bars.full_gy));
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      constexpr uint32_t idesc_kk = umma_idesc(128, 128, 0, 0);  // forward / dgrad: both operands K-major
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
      auto wgrad = [&](uint32_t d_tmem, uint32_t a_off, uint32_t b_off, bool first) {
#pragma unroll
        for (int j = 0; j < 8; ++j)  // 8 steps of 16 rows
          umma_f16(d_tmem, umma_desc(s_base + a_off + j * kMNmajKstep, kMNmajLbo, kSboBytes),
                   umma_desc(s_base + b_off + j * kMNmajKstep, kMNmajLbo, kSboBytes), idesc_mn,
                   (!first || j) ? 1u : 0u);
      };
      mbar_wait(smem_u32(&bars.w_full), 0);
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        const uint32_t par = it & 1u;
        // A1 of this tile landing implies that epilogue 2 of the previous tile is done with tmp2
        mbar_wait(smem_u32(&bars.full_a1), par);
        tc_fence_after();
        gemm_kk(t_tmp2, kBwSmemA1, kBwSmemWf);               // t = A1 Wa
        umma_commit(smem_u32(&bars.t_ready));
        mbar_wait(smem_u32(&bars.full_gy), par);
        tc_fence_after();
        gemm_kk(t_tmp1, kBwSmemGy, kBwSmemWb);               // g_a2 = g_y Wb^T
        umma_commit(smem_u32(&bars.tmp1_ready));
        mbar_wait(smem_u32(&bars.a2_ready), par);
        tc_fence_after();
        wgrad(t_dwb, kBwSmemA2, kBwSmemGy, it == 0);         // dWb += A2^T g_y
        umma_commit(smem_u32(&bars.free_gy));
        umma_commit(smem_u32(&bars.a2_free));
        mbar_wait(smem_u32(&bars.gt_ready), par);
        tc_fence_after();
        gemm_kk(t_tmp2, kBwSmemGt, kBwSmemWa);               // g_a1 = g_t Wa^T
        umma_commit(smem_u32(&bars.tmp2_ready));
        wgrad(t_dwa, kBwSmemA1, kBwSmemGt, it == 0);         // dWa += A1^T g_t
        umma_commit(smem_u32(&bars.free_a1));
        umma_commit(smem_u32(&bars.gt_free));
      }
      umma_commit(smem_u32(&bars.all_done));
    }
  } else if (warp >= 10) {
    // =========================== bias gradients ===========================
    // column sums of the g_y and g_t tiles straight from shared memory; each lane owns two columns of
    // slab `sl` for the whole kernel (two fp32 accumulators per tile kind), one flush at the end
    const uint32_t sl = warp - 10;
    const uint32_t chunk = lane >> 2, within = (lane & 3u) * 4u;
    float bb0 = 0.f, bb1 = 0.f, ba0 = 0.f, ba1 = 0.f;
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const uint32_t par = it & 1u;
      mbar_wait(smem_u32(&bars.full_gy), par);
      const uint8_t *gy = smem + kBwSmemGy + sl * CN_SLAB_BYTES + within;
#pragma unroll 8
      for (uint32_t r = 0; r < 128; ++r) {
        const uint32_t v = *reinterpret_cast<const uint32_t *>(gy + r * 128u + ((chunk ^ (r & 7u)) << 4));
        bb0 += bf16_lo(v);
        bb1 += bf16_hi(v);
      }
      mbar_arrive(smem_u32(&bars.free_gy));
      mbar_wait(smem_u32(&bars.gt_ready), par);
      const uint8_t *gt = smem + kBwSmemGt + sl * CN_SLAB_BYTES + within;
#pragma unroll 8
      for (uint32_t r = 0; r < 128; ++r) {
        const uint32_t v = *reinterpret_cast<const uint32_t *>(gt + r * 128u + ((chunk ^ (r & 7u)) << 4));
        ba0 += bf16_lo(v);
        ba1 += bf16_hi(v);
      }
      mbar_arrive(smem_u32(&bars.gt_free));
    }
    const int col = sl * 64 + lane * 2;
    atomicAdd(&g_bb[col], bb0);
    atomicAdd(&g_bb[col + 1], bb1);
    atomicAdd(&g_ba[col], ba0);
    atomicAdd(&g_ba[col + 1], ba1);
  } else {
    // =========================== epilogue: 8 warps, (row, column half) per thread ===========================
    const uint32_t e = warp - 2;
    const uint32_t half = e >> 2;
    const uint32_t quad = warp & 3u;
    const uint32_t row = quad * 32u + lane;
    const uint32_t lane_off = (quad * 32u) << 16;
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const uint32_t par = it & 1u;
      // ---- keep this half row of g_y in registers (needed again by epilogue 2), release the smem tile
      mbar_wait(smem_u32(&bars.full_gy), par);
      uint4 gy[8];
#pragma unroll
      for (int c = 0; c < 8; ++c)
        gy[c] = *reinterpret_cast<const uint4 *>(smem + kBwSmemGy + tile_chunk_off(row, half * 8 + c));
      mbar_arrive(smem_u32(&bars.free_gy));
      // ---- epilogue 0: A2 = relu(A1 Wa + ba) -> bf16 operand tile; its sign mask stays in registers
      uint32_t mask[2] = {0u, 0u};
      mbar_wait(smem_u32(&bars.t_ready), par);
      mbar_wait(smem_u32(&bars.a2_free), par ^ 1u);  // previous tile's wgrad-b is done with A2
      tc_fence_after();
      {
        uint32_t r[64];
        tmem_ld32(t_tmp2 + lane_off + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
        tmem_ld32(t_tmp2 + lane_off + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 b0 = *reinterpret_cast<const float4 *>(s_ba + half * 64 + q * 8);
          const float4 b1 = *reinterpret_cast<const float4 *>(s_ba + half * 64 + q * 8 + 4);
          float v[8];
          v[0] = __uint_as_float(r[q * 8 + 0]) + b0.x; v[1] = __uint_as_float(r[q * 8 + 1]) + b0.y;
          v[2] = __uint_as_float(r[q * 8 + 2]) + b0.z; v[3] = __uint_as_float(r[q * 8 + 3]) + b0.w;
          v[4] = __uint_as_float(r[q * 8 + 4]) + b1.x; v[5] = __uint_as_float(r[q * 8 + 5]) + b1.y;
          v[6] = __uint_as_float(r[q * 8 + 6]) + b1.z; v[7] = __uint_as_float(r[q * 8 + 7]) + b1.w;
          uint32_t ow[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            ow[i] = pack_bf16x2(fmaxf(v[2 * i], 0.f), fmaxf(v[2 * i + 1], 0.f));
            // the mask of the BF16 value (what the forward stored and the old backward tested)
            if (ow[i] & 0x0000ffffu) mask[q >> 2] |= 1u << ((q & 3) * 8 + 2 * i);
            if (ow[i] & 0xffff0000u) mask[q >> 2] |= 1u << ((q & 3) * 8 + 2 * i + 1);
          }
          *reinterpret_cast<uint4 *>(smem + kBwSmemA2 + tile_chunk_off(row, half * 8 + q)) =
              make_uint4(ow[0], ow[1], ow[2], ow[3]);
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(smem_u32(&bars.a2_ready));
      // ---- epilogue 1: g_t = g_a2 * [A2 > 0]
      mbar_wait(smem_u32(&bars.tmp1_ready), par);
      mbar_wait(smem_u32(&bars.gt_free), par ^ 1u);  // previous tile's wgrad-a / bias sums are done with GT
      tc_fence_after();
      {
        uint32_t r[64];
        tmem_ld32(t_tmp1 + lane_off + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
        tmem_ld32(t_tmp1 + lane_off + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const uint32_t m = mask[q >> 2] >> ((q & 3) * 8);
          uint32_t ow[4];
#pragma unroll
          for (int i = 0; i < 4; ++i)
            ow[i] = pack_bf16x2((m >> (2 * i)) & 1u ? __uint_as_float(r[q * 8 + 2 * i]) : 0.f,
                                (m >> (2 * i + 1)) & 1u ? __uint_as_float(r[q * 8 + 2 * i + 1]) : 0.f);
          *reinterpret_cast<uint4 *>(smem + kBwSmemGt + tile_chunk_off(row, half * 8 + q)) =
              make_uint4(ow[0], ow[1], ow[2], ow[3]);
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(smem_u32(&bars.gt_ready));
      // ---- epilogue 2: g_h = g_y + g_a1 * [A1 > 0]  -> global (one full 128-byte line per thread)
      mbar_wait(smem_u32(&bars.tmp2_ready), par);
      tc_fence_after();
      {
        uint32_t r[64];
        tmem_ld32(t_tmp2 + lane_off + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
        tmem_ld32(t_tmp2 + lane_off + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
        tmem_wait_ld();
        uint8_t *dst = gh_tiles + tile * (int64_t)CN_ACT_TILE_BYTES;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const uint4 m = *reinterpret_cast<const uint4 *>(smem + kBwSmemA1 + tile_chunk_off(row, half * 8 + q));
          const uint32_t mw[4] = {m.x, m.y, m.z, m.w};
          const uint32_t gw[4] = {gy[q].x, gy[q].y, gy[q].z, gy[q].w};
          uint32_t ow[4];
#pragma unroll
          for (int i = 0; i < 4; ++i)
            ow[i] = pack_bf16x2(
                bf16_lo(gw[i]) + (bf16_lo(mw[i]) > 0.f ? __uint_as_float(r[q * 8 + 2 * i]) : 0.f),
                bf16_hi(gw[i]) + (bf16_hi(mw[i]) > 0.f ? __uint_as_float(r[q * 8 + 2 * i + 1]) : 0.f));
          *reinterpret_cast<uint4 *>(dst + tile_chunk_off(row, half * 8 + q)) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&bars.free_a1));
    }
    // ---- flush the weight gradients held in TMEM (each thread: its input unit, 64 output units)
    mbar_wait(smem_u32(&bars.all_done), 0);
    tc_fence_after();
#pragma unroll 1
    for (int mtx = 0; mtx < 2; ++mtx) {
      float *gdst = (mtx == 0 ? g_wa : g_wb) + row * 128 + half * 64;  // accumulator row = input unit
      const uint32_t tsrc = (mtx == 0 ? t_dwa : t_dwb) + lane_off + half * 64;
#pragma unroll 1
      for (int c2 = 0; c2 < 2; ++c2) {
        uint32_t r[32];
        tmem_ld32(tsrc + c2 * 32, r);
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 8; ++q)
          red_add_f4(gdst + c2 * 32 + q * 4, __uint_as_float(r[q * 4 + 0]), __uint_as_float(r[q * 4 + 1]),
                     __uint_as_float(r[q * 4 + 2]), __uint_as_float(r[q * 4 + 3]));
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
__global__ void __launch_bounds__(128)
readout_bwd_kernel(const float *__restrict__ wr, const float *__restrict__ d_y, const uint8_t *__restrict__ acts,
                   int64_t act_stride_bytes, int64_t n_tiles, uint8_t *__restrict__ g_tiles,
                   float *__restrict__ g_wr, float *__restrict__ g_br) {
  __shared__ float s_wr[512];
  __shared__ float s_dy[128 * 4];
  __shared__ __align__(16) uint8_t s_a[CN_ACT_TILE_BYTES];
  const int tid = threadIdx.x;
  for (int i = tid; i < 512; i += 128) s_wr[i] = wr[i];
  float acc[4] = {0.f, 0.f, 0.f, 0.f};  // dWr[tid][0..3]
  float accb = 0.f;                     // dbr[tid] for tid < 4
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    __syncthreads();
    const float4 dy = reinterpret_cast<const float4 *>(d_y)[tile * 128 + tid];
    reinterpret_cast<float4 *>(s_dy)[tid] = dy;
    const uint8_t *src = acts + tile * act_stride_bytes;
    uint8_t *dst = g_tiles + tile * (int64_t)CN_ACT_TILE_BYTES;
    for (int c = 0; c < 16; ++c) {
      const uint32_t off = tile_chunk_off(tid, c);
      const uint4 a = ldg_nc_u4(src + off);
      *reinterpret_cast<uint4 *>(s_a + off) = a;
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
      uint32_t ow[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k0 = c * 8 + 2 * i;
        const float g0 = dy.x * s_wr[k0 * 4 + 0] + dy.y * s_wr[k0 * 4 + 1] + dy.z * s_wr[k0 * 4 + 2] +
                         dy.w * s_wr[k0 * 4 + 3];
        const float g1 = dy.x * s_wr[k0 * 4 + 4] + dy.y * s_wr[k0 * 4 + 5] + dy.z * s_wr[k0 * 4 + 6] +
                         dy.w * s_wr[k0 * 4 + 7];
        ow[i] = pack_bf16x2(bf16_lo(aw[i]) > 0.f ? g0 : 0.f, bf16_hi(aw[i]) > 0.f ? g1 : 0.f);
      }
      *reinterpret_cast<uint4 *>(dst + off) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
    }
    __syncthreads();
    // thread k accumulates dWr[k][:] over the 128 rows of the tile
    for (int r = 0; r < 128; ++r) {
      const float a = __bfloat162float(*reinterpret_cast<const __nv_bfloat16 *>(s_a + tile_off(r, tid)));
      acc[0] += a * s_dy[r * 4 + 0];
      acc[1] += a * s_dy[r * 4 + 1];
      acc[2] += a * s_dy[r * 4 + 2];
      acc[3] += a * s_dy[r * 4 + 3];
    }
    if (tid < 4)
      for (int r = 0; r < 128; ++r) accb += s_dy[r * 4 + tid];
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) atomicAdd(&g_wr[tid * 4 + j], acc[j]);
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

extern "C" int cn_mlp_block_bwd(const void *packed, int32_t blk, const void *acts, int32_t act_stride_tiles,
                                int32_t act_index, int64_t n_tiles, const void *gy_tiles, void *gh_tiles, float *grad,
                                void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && acts && gy_tiles && gh_tiles && grad, "cn_mlp_block_bwd: null pointer");
  CN_REQUIRE(blk >= 0 && blk < 6, "cn_mlp_block_bwd: blk %d", blk);
  CN_REQUIRE(act_stride_tiles >= 1 && act_index >= 0 && act_index < act_stride_tiles,
             "cn_mlp_block_bwd: act_stride_tiles=%d act_index=%d", act_stride_tiles, act_index);
  CN_REQUIRE(n_tiles >= 0, "cn_mlp_block_bwd: n_tiles %lld", (long long)n_tiles);
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
  const unsigned grid = (unsigned)(n_tiles < kNumSMs ? n_tiles : kNumSMs);
  mlp_block_bwd_kernel<<<grid, kBwdThreads, kBwdSmemBytes, (cudaStream_t)stream>>>(
      pk + pack_bwd_block(blk, 0), pk + pack_bwd_block(blk, 1), pk + pack_fwd_block(blk, 0),
      reinterpret_cast<const float *>(pk + kPackTables) + (blk < 3 ? kTabFeature : kTabFusion) + (2 * (blk % 3) + 1) * 128,
      (const uint8_t *)acts + (int64_t)act_index * CN_ACT_TILE_BYTES, (int64_t)act_stride_tiles * CN_ACT_TILE_BYTES,
      n_tiles, (const uint8_t *)gy_tiles, (uint8_t *)gh_tiles, grad + param_block_w(blk, 0),
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
  CN_REQUIRE((((uintptr_t)d_y | (uintptr_t)acts | (uintptr_t)g_tiles) & 15) == 0, "cn_mlp_readout_bwd: alignment");
  if (n_tiles == 0) return 0;
  const unsigned grid = (unsigned)(n_tiles < 4 * kNumSMs ? n_tiles : 4 * kNumSMs);
  readout_bwd_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(
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
  mean_bwd_kernel<<<kNumSMs * 8, 256, 0, (cudaStream_t)stream>>>((const uint4 *)gm_tiles, b, nv, tiles_per_view,
                                                                 (uint4 *)gh3_tiles);
  return check_launch("cn_mlp_mean_bwd");
}
