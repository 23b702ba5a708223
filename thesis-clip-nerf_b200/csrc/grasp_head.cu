// GraspReadout (delta_ngf/layers.py:8-42 = lmvnerf/layers.py; relu variant grasp_mvnerf/layers.py:8-31) on tcgen05 / TMEM:
// the affordance head of the pose-query models, forward and the backward to its inputs (the gradient the test-time
// pose optimiser takes, grasp_optimizer.py:158-184).
//
//   per query point (row):  a_s = act(tap_s W_s + b_s), s = 1..4 (128 -> 64);  c = act([a_1 a_2 a_3 a_4] W_c + b_c) (256 -> 64)
//   per pose (42 points):   v = flatten(c) (2688);  r0 = act(v) W_0 + b_0 (2688 -> 128);  o1 = v W_sc + act(r0) W_1 + b_1 (64)
//                           o2 = o1 + act(act(o1) M_0 + m_0) M_1 + m_1;   y = relu(o2) w_o + b_o
//
// Four kernels.  "point" kernels work on the 128-row tiles of the fusion stage (rows = (scene, pose, offset)); "pose"
// kernels on tiles of 128 poses whose 2688-wide input is streamed as 42 K-chunks of 64:
//   grasp_point_fwd_kernel  taps (bf16 tiles written by the fusion stage's epilogues = the A operands) -> 4 + 1 GEMMs -> V slabs: per (pose tile, offset j) a
//                           [128 poses x 64] bf16 slab of v and one of act(v), i.e. the K-chunks of the pose kernel's A operand,
//                           written straight from the epilogue registers (the 'b np n5 d -> b np (n5 d)' rearrange is the store address)
//   grasp_pose_fwd_kernel   TMA ring over the 42 chunks (V pair + W_0 chunk + W_sc chunk), two accumulators, then the
//                           three small GEMMs of the tail with the epilogue warps producing each next operand in shared memory
//   grasp_pose_bwd_kernel   tail backward (thread per pose + 3 GEMMs), then 42 chunks of d v = d_sc W_sc^T + (d_r0 W_0^T) act'(v),
//                           chained through act'(c): double-buffered TMEM accumulators, epilogue writes the d(pre-activation of c) slabs
//   grasp_point_bwd_kernel  d cat = d_c W_c^T (N = 256), d a_s = d cat_s act'(a_s) in place over the saved a_s tile, d tap_s = d a_s W_s^T
//                           -> bf16 gradient tiles that enter the gradient stream of cn_mlp_block_bwd (cn_tile_add_tiles)
// act' is taken from the activation's OUTPUT (elu: y > 0 ? 1 : y + 1; relu: y > 0), so only outputs are saved, as bf16
// operand tiles.  Weights: bf16 K-major images of every matrix in both orientations + one fp32 bias table, built once
// per weight version by cn_grasp_head_pack from the Keras-order flat parameter vector.
#include "common.cuh"
#include <mutex>

namespace cn {

// ---- flat fp32 parameters, Keras variable order (oracle grasp_readout_param_shapes), kernel[in][out]
constexpr int kHpDs = 0, kHpDsStride = 128 * 64 + 64;          // activation_downscale_{1..4}: kernel, bias
constexpr int kHpComb = 4 * kHpDsStride, kHpCombB = kHpComb + 256 * 64;
constexpr int kHpL0 = kHpCombB + 64, kHpL0B = kHpL0 + 2688 * 128;
constexpr int kHpL1 = kHpL0B + 128, kHpL1B = kHpL1 + 128 * 64;
constexpr int kHpSc = kHpL1B + 64;
constexpr int kHpM0 = kHpSc + 2688 * 64, kHpM0B = kHpM0 + 64 * 64;
constexpr int kHpM1 = kHpM0B + 64, kHpM1B = kHpM1 + 64 * 64;
constexpr int kHpOut = kHpM1B + 64, kHpOutB = kHpOut + 64;
static_assert(kHpOutB + 1 == CN_HEAD_PARAMS, "GraspReadout parameter count");

// ---- packed images (byte offsets).  image(N, K): K/64 slabs of [N rows x 64 k] bf16, 128-byte swizzle
constexpr uint32_t img(uint32_t N, uint32_t K) { return (K / 64) * N * 128; }
constexpr uint32_t kHfDs = 0;                                   // forward: B[n = out][k = in]
constexpr uint32_t kHfComb = kHfDs + 4 * img(64, 128);
constexpr uint32_t kHfL1 = kHfComb + img(64, 256);
constexpr uint32_t kHfM0 = kHfL1 + img(64, 128);
constexpr uint32_t kHfM1 = kHfM0 + img(64, 64);
constexpr uint32_t kHfL0 = kHfM1 + img(64, 64);
constexpr uint32_t kHfSc = kHfL0 + img(128, 2688);
constexpr uint32_t kHbDs = kHfSc + img(64, 2688);               // backward: B[n = in][k = out]
constexpr uint32_t kHbComb = kHbDs + 4 * img(128, 64);
constexpr uint32_t kHbL1 = kHbComb + img(256, 64);
constexpr uint32_t kHbM0 = kHbL1 + img(128, 64);
constexpr uint32_t kHbM1 = kHbM0 + img(64, 64);
constexpr uint32_t kHbL0 = kHbM1 + img(64, 64);
constexpr uint32_t kHbSc = kHbL0 + img(2688, 128);
constexpr uint32_t kHeadTab = kHbSc + img(2688, 64);
// fp32 table (float offsets)
constexpr int kTbDs = 0, kTbComb = 256, kTbL0 = 320, kTbL1 = 448, kTbM0 = 512, kTbM1 = 576, kTbOut = 640, kTbOutB = 704;
constexpr int kHeadTabFloats = 768;
static_assert(kHeadTab + kHeadTabFloats * 4 == CN_HEAD_PACK_BYTES, "head pack size");
static_assert(kHfComb == 65536 && kHfL1 == 98304 && kHfL0 == 131072, "head pack layout");

constexpr int kN5 = 42;                                          // query offsets per pose (6 bases x 7 steps)
constexpr uint32_t kVPair = 2 * CN_SLAB_BYTES;                   // v slab | act(v) slab
constexpr uint32_t kSavedTile = 5 * CN_SLAB_BYTES;               // act(r0) (2 slabs) | act(o1) | act(r2) | relu(o2)
constexpr uint32_t kAcatTile = 4 * CN_SLAB_BYTES;                // a_1 .. a_4

// elu's negative branch as ex2.approx(x log2 e) - 1: absolute error ~1e-7 (the cancellation near 0 costs relative, not
// absolute, accuracy), far inside the bf16 rounding of the stored output; expm1f made the forward epilogues 2x longer
__device__ __forceinline__ float act_fwd(float x, int elu) { return x > 0.f ? x : (elu ? __expf(x) - 1.f : 0.f); }
// derivative of the activation from its OUTPUT y
__device__ __forceinline__ float act_der(float y, int elu) { return y > 0.f ? 1.f : (elu ? y + 1.f : 0.f); }

// ------------------------------------------------------------------------------------------------ weight pack
struct PackEntry {
  int src, n_stride, k_stride, N, K;
  uint32_t dst;
};
struct PackTable {
  PackEntry e[28];
  int n;
};
__global__ void __launch_bounds__(256) grasp_head_pack_kernel(const float *__restrict__ params, uint8_t *__restrict__ pack,
                                                              PackTable tab) {
  if ((int)blockIdx.y == tab.n) {
    // fp32 biases + the output layer
    float *t = reinterpret_cast<float *>(pack + kHeadTab);
    for (int i = blockIdx.x * 256 + threadIdx.x; i < kHeadTabFloats; i += gridDim.x * 256) {
      float v = 0.f;
      if (i < 256) v = params[kHpDs + (i >> 6) * kHpDsStride + 128 * 64 + (i & 63)];
      else if (i < 320) v = params[kHpCombB + i - 256];
      else if (i < 448) v = params[kHpL0B + i - 320];
      else if (i < 512) v = params[kHpL1B + i - 448];
      else if (i < 576) v = params[kHpM0B + i - 512];
      else if (i < 640) v = params[kHpM1B + i - 576];
      else if (i < 704) v = params[kHpOut + i - 640];
      else if (i == 704) v = params[kHpOutB];
      t[i] = v;
    }
    return;
  }
  const PackEntry E = tab.e[blockIdx.y];
  const int n_chunks = E.N * (E.K / 8);
  for (int q = blockIdx.x * 256 + threadIdx.x; q < n_chunks; q += gridDim.x * 256) {
    const int n = q / (E.K / 8), kc = q % (E.K / 8);
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = params[E.src + (int64_t)n * E.n_stride + (int64_t)(kc * 8 + i) * E.k_stride];
    uint8_t *dst = pack + E.dst + (uint32_t)(kc >> 3) * (uint32_t)E.N * 128u + (uint32_t)n * 128u +
                   ((((uint32_t)kc & 7u) ^ ((uint32_t)n & 7u)) << 4);
    *reinterpret_cast<uint4 *>(dst) =
        make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  }
}

// rows are (scene, pose, offset): row R -> scene R / n_pad, i = R % n_pad, valid i < np * 42, pose = scene * np + i / 42
struct HeadGeom {
  int64_t n_tiles;      // row tiles (b * n_pad / 128)
  int n_pad, np;        // rows per scene (multiple of 128), poses per scene
  int n_poses;          // b * np
  int elu;              // 1: elu head (DeltaNGF / LanguageNeRF), 0: relu head (GraspMVNeRF)
};
__device__ __forceinline__ bool row_pose(const HeadGeom &g, int64_t R, int &pose, int &j) {
  const int scene = (int)(R / g.n_pad), i = (int)(R % g.n_pad);
  pose = scene * g.np + i / kN5;
  j = i % kN5;
  return i < g.np * kN5;
}

// ------------------------------------------------------------------------------------------------ point forward
constexpr int kPtThreads = 256;
constexpr uint32_t kPfSmemW = 0;                       // pack [0, 98304): W_1..4 (4 x 16 KiB), W_c (32 KiB)
constexpr uint32_t kPfSmemA = 98304;                   // 2 x 32 KiB tap operand tiles (TMA ring)
constexpr uint32_t kPfSmemCat = kPfSmemA + 2 * CN_ACT_TILE_BYTES;   // [a_1 .. a_4]: 4 slabs
constexpr uint32_t kPfSmemBytes = kPfSmemCat + kAcatTile + 1024;

// taps: bf16 activation-tile images [4][n_tiles][32 KiB] (cn_mlp_fusion_fwd_tap_tiles): a tap tile IS the A operand
// of its GEMM, one bulk copy each through a two-slot ring.  Thread 0 feeds the ring and issues the GEMMs; the loads
// of the next tile's first two taps fly under this tile's epilogues.
__global__ void __launch_bounds__(kPtThreads, 1)
grasp_point_fwd_kernel(const uint8_t *__restrict__ taps, int64_t tap_stride, const uint8_t *__restrict__ pack, HeadGeom g,
                       uint8_t *__restrict__ vslabs, uint8_t *__restrict__ acat) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_w, bar_full[2], bar_mma[2], bar_taps, bar_c;
  __shared__ uint32_t tmem_slot;
  __shared__ float s_bias[320];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t s_base = smem_u32(smem);
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
  for (int i = tid; i < 320; i += kPtThreads) s_bias[i] = reinterpret_cast<const float *>(pack + kHeadTab)[i];
  auto load_tap = [&](int64_t tile, int s) {           // thread 0 only
    const uint32_t buf = (uint32_t)s & 1u;
    mbar_expect_tx(smem_u32(&bar_full[buf]), CN_ACT_TILE_BYTES);
    bulk_g2s(s_base + kPfSmemA + buf * CN_ACT_TILE_BYTES, taps + s * tap_stride + tile * (int64_t)CN_ACT_TILE_BYTES,
             CN_ACT_TILE_BYTES, smem_u32(&bar_full[buf]));
  };
  if (tid == 0) {
    mbar_init(smem_u32(&bar_w), 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&bar_full[i]), 1);
      mbar_init(smem_u32(&bar_mma[i]), 1);
    }
    mbar_init(smem_u32(&bar_taps), 1);
    mbar_init(smem_u32(&bar_c), 1);
    fence_barrier_init();
    mbar_expect_tx(smem_u32(&bar_w), 98304);
    bulk_g2s(s_base + kPfSmemW, pack, 98304, smem_u32(&bar_w));
    if ((int64_t)blockIdx.x < g.n_tiles) {
      load_tap(blockIdx.x, 0);
      load_tap(blockIdx.x, 1);
    }
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t quad = warp & 3u, half = warp >> 2, row = quad * 32u + lane;
  const uint32_t lane_off = (quad * 32u) << 16;
  constexpr uint32_t idesc64 = umma_idesc(128, 64, 0, 0);

  uint32_t it = 0;
  for (int64_t tile = blockIdx.x; tile < g.n_tiles; tile += gridDim.x, ++it) {
    if (tid == 0) {
      if (it == 0) mbar_wait(smem_u32(&bar_w), 0);
      const int64_t next = tile + gridDim.x;
      for (int s = 0; s < 4; ++s) {
        const uint32_t buf = (uint32_t)s & 1u, ph = (2u * it + ((uint32_t)s >> 1)) & 1u;   // 2 fills per slot and tile
        mbar_wait(smem_u32(&bar_full[buf]), ph);
        tc_fence_after();
        const uint32_t a = s_base + kPfSmemA + buf * CN_ACT_TILE_BYTES, w = s_base + kPfSmemW + s * img(64, 128);
#pragma unroll
        for (int sl = 0; sl < 2; ++sl)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16(tmem + s * 64, umma_desc(a + sl * CN_SLAB_BYTES + k * kKmajKstep, kKmajLbo, kSboBytes),
                     umma_desc(w + sl * 8192 + k * kKmajKstep, kKmajLbo, kSboBytes), idesc64, (sl | k) ? 1u : 0u);
        umma_commit(smem_u32(&bar_mma[buf]));
        if (s == 3) umma_commit(smem_u32(&bar_taps));
        // refill the slot with the tap two ahead once this GEMM has read it
        if (s < 2 || next < g.n_tiles) {
          mbar_wait(smem_u32(&bar_mma[buf]), ph);
          if (s < 2) load_tap(tile, s + 2);
          else load_tap(next, s - 2);
        }
      }
    }
    mbar_wait(smem_u32(&bar_taps), it & 1u);
    tc_fence_after();
    // ---- a_s = act(tap_s W_s + b_s) -> slab s of the [128 x 256] operand of the combined downscale
    for (int ss = 0; ss < 2; ++ss) {
      const int s = 2 * (int)half + ss;
      uint32_t v[64];
      tmem_ld32(tmem + lane_off + s * 64, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tmem_ld32(tmem + lane_off + s * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tmem_wait_ld();
      uint8_t *dst = smem + kPfSmemCat + s * CN_SLAB_BYTES + row * 128u;
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8) {
        float a[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = act_fwd(__uint_as_float(v[c8 * 8 + i]) + s_bias[s * 64 + c8 * 8 + i], g.elu);
        *reinterpret_cast<uint4 *>(dst + (((uint32_t)c8 ^ (row & 7u)) << 4)) =
            make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
      }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t a = s_base + kPfSmemCat, w = s_base + kPfSmemW + kHfComb;
#pragma unroll
      for (int sl = 0; sl < 4; ++sl)
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_f16(tmem + 256, umma_desc(a + sl * CN_SLAB_BYTES + k * kKmajKstep, kKmajLbo, kSboBytes),
                   umma_desc(w + sl * 8192 + k * kKmajKstep, kKmajLbo, kSboBytes), idesc64, (sl | k) ? 1u : 0u);
      umma_commit(smem_u32(&bar_c));
      if (acat) {                                     // saved for the backward: the operand tile as it stands
        bulk_s2g(acat + tile * (int64_t)kAcatTile, a, kAcatTile);
        bulk_commit();
      }
    }
    mbar_wait(smem_u32(&bar_c), it & 1u);
    tc_fence_after();
    // ---- c = act(. + b_c): this thread's 32 columns go to row (pose % 128) of the V slabs of (pose tile, offset j)
    {
      uint32_t v[32];
      tmem_ld32(tmem + lane_off + 256 + half * 32, v);
      tmem_wait_ld();
      int pose, j;
      if (row_pose(g, tile * 128 + row, pose, j)) {
        uint8_t *dst = vslabs + ((int64_t)(pose >> 7) * kN5 + j) * kVPair + (uint32_t)(pose & 127) * 128u;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float c[8], e[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            c[i] = act_fwd(__uint_as_float(v[q * 8 + i]) + s_bias[256 + half * 32 + q * 8 + i], g.elu);
            e[i] = act_fwd(c[i], g.elu);              // ResNetMLPBlock's prelu_0 on an already activated input
          }
          const uint32_t off = ((half * 4u + (uint32_t)q) ^ ((uint32_t)pose & 7u)) << 4;
          *reinterpret_cast<uint4 *>(dst + off) =
              make_uint4(pack_bf16x2(c[0], c[1]), pack_bf16x2(c[2], c[3]), pack_bf16x2(c[4], c[5]), pack_bf16x2(c[6], c[7]));
          *reinterpret_cast<uint4 *>(dst + CN_SLAB_BYTES + off) =
              make_uint4(pack_bf16x2(e[0], e[1]), pack_bf16x2(e[2], e[3]), pack_bf16x2(e[4], e[5]), pack_bf16x2(e[6], e[7]));
        }
      }
    }
    tc_fence_before();
    if (tid == 0 && acat) bulk_wait_read0();          // the store has read the operand tile: the next tile may overwrite it
    __syncthreads();
  }
  if (tid == 0) bulk_wait0();
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ pose forward
constexpr int kPoseThreads = 192;                      // warp 0: TMA producer, warp 1: MMA issuer, warps 2-5: epilogue
constexpr uint32_t kPsStage = kVPair + img(128, 64) + img(64, 64);   // V pair | W_0 chunk | W_sc chunk = 56 KiB
constexpr int kPsStages = 3;
constexpr uint32_t kPsSmemTailW = kPsStages * kPsStage;              // W_1 | M_0 | M_1 (pack [kHfL1, kHfL0))
constexpr uint32_t kPsSmemBytes = kPsSmemTailW + (kHfL0 - kHfL1) + 1024;

__global__ void __launch_bounds__(kPoseThreads, 1)
grasp_pose_fwd_kernel(const uint8_t *__restrict__ vslabs, const uint8_t *__restrict__ pack, HeadGeom g,
                      uint8_t *__restrict__ saved, float *__restrict__ y) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[kPsStages], bar_empty[kPsStages], bar_w, bar_main, bar_tail;
  __shared__ uint32_t tmem_slot;
  __shared__ float s_tab[kHeadTabFloats];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t s_base = smem_u32(smem);
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
  const int64_t ptile = blockIdx.x;
  for (int i = tid; i < kHeadTabFloats; i += kPoseThreads) s_tab[i] = reinterpret_cast<const float *>(pack + kHeadTab)[i];
  if (tid == 0) {
    for (int i = 0; i < kPsStages; ++i) {
      mbar_init(smem_u32(&bar_full[i]), 1);
      mbar_init(smem_u32(&bar_empty[i]), 1);
    }
    mbar_init(smem_u32(&bar_w), 1);
    mbar_init(smem_u32(&bar_main), 1);
    mbar_init(smem_u32(&bar_tail), 1);
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
  constexpr uint32_t kOffW0 = kVPair, kOffSc = kVPair + img(128, 64);
  constexpr uint32_t idesc128 = umma_idesc(128, 128, 0, 0), idesc64 = umma_idesc(128, 64, 0, 0);

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(smem_u32(&bar_w), kHfL0 - kHfL1);
      bulk_g2s(s_base + kPsSmemTailW, pack + kHfL1, kHfL0 - kHfL1, smem_u32(&bar_w));
      for (int j = 0; j < kN5; ++j) {
        const uint32_t st = (uint32_t)j % kPsStages, dst = s_base + st * kPsStage;
        mbar_wait(smem_u32(&bar_empty[st]), (((uint32_t)j / kPsStages) & 1u) ^ 1u);
        mbar_expect_tx(smem_u32(&bar_full[st]), kPsStage);
        bulk_g2s(dst, vslabs + (ptile * kN5 + j) * (int64_t)kVPair, kVPair, smem_u32(&bar_full[st]));
        bulk_g2s(dst + kOffW0, pack + kHfL0 + (uint32_t)j * img(128, 64), img(128, 64), smem_u32(&bar_full[st]));
        bulk_g2s(dst + kOffSc, pack + kHfSc + (uint32_t)j * img(64, 64), img(64, 64), smem_u32(&bar_full[st]));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      for (int j = 0; j < kN5; ++j) {
        const uint32_t st = (uint32_t)j % kPsStages, a = s_base + st * kPsStage;
        mbar_wait(smem_u32(&bar_full[st]), ((uint32_t)j / kPsStages) & 1u);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k)      // r0 += act(v)_j W_0[j]
          umma_f16(tmem, umma_desc(a + CN_SLAB_BYTES + k * kKmajKstep, kKmajLbo, kSboBytes),
                   umma_desc(a + kOffW0 + k * kKmajKstep, kKmajLbo, kSboBytes), idesc128, (j | k) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 4; ++k)      // shortcut += v_j W_sc[j]
          umma_f16(tmem + 128, umma_desc(a + k * kKmajKstep, kKmajLbo, kSboBytes),
                   umma_desc(a + kOffSc + k * kKmajKstep, kKmajLbo, kSboBytes), idesc64, (j | k) ? 1u : 0u);
        umma_commit(smem_u32(&bar_empty[st]));
      }
      umma_commit(smem_u32(&bar_main));
    }
  } else {
    // ---- tail: thread = pose.  Each step turns an accumulator into the next GEMM's operand tile in shared memory
    // (over ring stage 0: the ring is drained), saves it for the backward, and the leader issues the GEMM.
    const uint32_t quad = warp & 3u, row = quad * 32u + lane, lane_off = (quad * 32u) << 16;
    const bool leader = warp == 2 && lane == 0;
    const int64_t pose = ptile * 128 + row;
    uint8_t *A = smem;                                   // operand tile of the tail GEMMs
    uint8_t *S = saved + ptile * (int64_t)kSavedTile + row * 128u;
    uint32_t pht = 0;
    auto chunk_out = [&](const float (&a)[8], uint32_t slab_a, uint32_t slab_s, uint32_t c8) {
      const uint4 o = make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
      const uint32_t off = (c8 ^ (row & 7u)) << 4;
      *reinterpret_cast<uint4 *>(A + slab_a * CN_SLAB_BYTES + row * 128u + off) = o;
      *reinterpret_cast<uint4 *>(S + slab_s * CN_SLAB_BYTES + off) = o;
    };
    auto gemm = [&](uint32_t d_col, uint32_t w_off, int k_slabs) {   // D[128 x 64] = A (k_slabs slabs) . W
      fence_proxy_async_smem();
      tc_fence_before();
      named_bar_sync(1, 128);
      if (leader) {
        tc_fence_after();
        for (int sl = 0; sl < k_slabs; ++sl)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16(tmem + d_col, umma_desc(s_base + sl * CN_SLAB_BYTES + k * kKmajKstep, kKmajLbo, kSboBytes),
                     umma_desc(s_base + kPsSmemTailW + w_off + sl * 8192 + k * kKmajKstep, kKmajLbo, kSboBytes), idesc64,
                     (sl | k) ? 1u : 0u);
        umma_commit(smem_u32(&bar_tail));
      }
      mbar_wait(smem_u32(&bar_tail), pht & 1u);
      ++pht;
      tc_fence_after();
    };
    mbar_wait(smem_u32(&bar_main), 0);
    mbar_wait(smem_u32(&bar_w), 0);
    tc_fence_after();
    // T1: act(r0 + b_0) (128)
#pragma unroll 1
    for (int c4 = 0; c4 < 4; ++c4) {
      uint32_t v[32];
      tmem_ld32(tmem + lane_off + c4 * 32, v);
      tmem_wait_ld();
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float a[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = act_fwd(__uint_as_float(v[q * 8 + i]) + s_tab[kTbL0 + c4 * 32 + q * 8 + i], g.elu);
        chunk_out(a, (uint32_t)c4 >> 1, (uint32_t)c4 >> 1, (uint32_t)(c4 & 1) * 4 + q);
      }
    }
    gemm(192, 0, 2);                                       // act(r0) W_1
    // T2: o1 = shortcut + . + b_1 (kept in registers), act(o1)
    float o1[64];
    {
      uint32_t v[64], w[64];
      tmem_ld32(tmem + lane_off + 128, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tmem_ld32(tmem + lane_off + 160, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tmem_ld32(tmem + lane_off + 192, *reinterpret_cast<uint32_t(*)[32]>(&w[0]));
      tmem_ld32(tmem + lane_off + 224, *reinterpret_cast<uint32_t(*)[32]>(&w[32]));
      tmem_wait_ld();
#pragma unroll
      for (int i = 0; i < 64; ++i) o1[i] = __uint_as_float(v[i]) + __uint_as_float(w[i]) + s_tab[kTbL1 + i];
    }
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) {
      float a[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = act_fwd(o1[c8 * 8 + i], g.elu);
      chunk_out(a, 0, 2, c8);
    }
    gemm(256, img(64, 128), 1);                            // act(o1) M_0
    // T3: act(. + m_0)
    {
      uint32_t v[64];
      tmem_ld32(tmem + lane_off + 256, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tmem_ld32(tmem + lane_off + 288, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tmem_wait_ld();
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8) {
        float a[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = act_fwd(__uint_as_float(v[c8 * 8 + i]) + s_tab[kTbM0 + c8 * 8 + i], g.elu);
        chunk_out(a, 0, 3, c8);
      }
    }
    gemm(320, img(64, 128) + img(64, 64), 1);              // . M_1
    // T4: o2 = o1 + . + m_1;  y = relu(o2) w_o + b_o
    {
      uint32_t v[64];
      tmem_ld32(tmem + lane_off + 320, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tmem_ld32(tmem + lane_off + 352, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tmem_wait_ld();
      float acc = s_tab[kTbOutB];
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8) {
        float a[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          a[i] = fmaxf(o1[c8 * 8 + i] + __uint_as_float(v[c8 * 8 + i]) + s_tab[kTbM1 + c8 * 8 + i], 0.f);
          acc += a[i] * s_tab[kTbOut + c8 * 8 + i];
        }
        *reinterpret_cast<uint4 *>(S + 4 * CN_SLAB_BYTES + (((uint32_t)c8 ^ (row & 7u)) << 4)) =
            make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
      }
      if (pose < g.n_poses) y[pose] = acc;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ pose backward
constexpr int kPbStages = 2;
constexpr uint32_t kPbStage = kVPair + img(64, 128) + img(64, 64);   // V pair | W_0^T chunk (2 k-slabs) | W_sc^T chunk
constexpr uint32_t kPbSmemW = kPbStages * kPbStage;                   // W_1^T | M_0^T | M_1^T (pack [kHbL1, kHbL0))
constexpr uint32_t kPbSmemG0 = kPbSmemW + (kHbL0 - kHbL1);            // d r0 operand tile [128 x 128]
constexpr uint32_t kPbSmemGs = kPbSmemG0 + CN_ACT_TILE_BYTES;         // d o1 operand tile [128 x 64] (and tail scratch)
constexpr uint32_t kPbSmemBytes = kPbSmemGs + CN_SLAB_BYTES + 1024;

__global__ void __launch_bounds__(kPoseThreads, 1)
grasp_pose_bwd_kernel(const float *__restrict__ gy, const uint8_t *__restrict__ vslabs, const uint8_t *__restrict__ saved,
                      const uint8_t *__restrict__ pack, HeadGeom g, uint8_t *__restrict__ dpc) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[kPbStages], bar_empty[kPbStages], bar_w, bar_tail, bar_g, bar_acc_full[2],
      bar_acc_free[2];
  __shared__ uint32_t tmem_slot;
  __shared__ float s_wo[64];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t s_base = smem_u32(smem);
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
  const int64_t ptile = blockIdx.x;
  if (tid < 64) s_wo[tid] = reinterpret_cast<const float *>(pack + kHeadTab)[kTbOut + tid];
  if (tid == 0) {
    for (int i = 0; i < kPbStages; ++i) {
      mbar_init(smem_u32(&bar_full[i]), 1);
      mbar_init(smem_u32(&bar_empty[i]), 1 + 4);        // MMA commit + the 4 epilogue warps (they read the V slabs)
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&bar_acc_full[i]), 1);
      mbar_init(smem_u32(&bar_acc_free[i]), 4);
    }
    mbar_init(smem_u32(&bar_w), 1);
    mbar_init(smem_u32(&bar_tail), 1);
    mbar_init(smem_u32(&bar_g), 4);
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
  constexpr uint32_t kOffW0 = kVPair, kOffSc = kVPair + img(64, 128);
  constexpr uint32_t idesc128 = umma_idesc(128, 128, 0, 0), idesc64 = umma_idesc(128, 64, 0, 0);

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(smem_u32(&bar_w), kHbL0 - kHbL1);
      bulk_g2s(s_base + kPbSmemW, pack + kHbL1, kHbL0 - kHbL1, smem_u32(&bar_w));
      for (int j = 0; j < kN5; ++j) {
        const uint32_t st = (uint32_t)j % kPbStages, dst = s_base + st * kPbStage;
        mbar_wait(smem_u32(&bar_empty[st]), (((uint32_t)j / kPbStages) & 1u) ^ 1u);
        mbar_expect_tx(smem_u32(&bar_full[st]), kPbStage);
        bulk_g2s(dst, vslabs + (ptile * kN5 + j) * (int64_t)kVPair, kVPair, smem_u32(&bar_full[st]));
        // rows [64 j, 64 j + 64) of the [2688 x 128] image: one 8 KiB piece of each of its two k-slabs
        bulk_g2s(dst + kOffW0, pack + kHbL0 + (uint32_t)j * 8192u, 8192, smem_u32(&bar_full[st]));
        bulk_g2s(dst + kOffW0 + 8192, pack + kHbL0 + 2688u * 128u + (uint32_t)j * 8192u, 8192, smem_u32(&bar_full[st]));
        bulk_g2s(dst + kOffSc, pack + kHbSc + (uint32_t)j * 8192u, 8192, smem_u32(&bar_full[st]));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      mbar_wait(smem_u32(&bar_g), 0);                   // d r0 / d o1 operand tiles are in shared memory
      for (int j = 0; j < kN5; ++j) {
        const uint32_t st = (uint32_t)j % kPbStages, a = s_base + st * kPbStage, acc = 256u + ((uint32_t)j & 1u) * 128u;
        mbar_wait(smem_u32(&bar_full[st]), ((uint32_t)j / kPbStages) & 1u);
        mbar_wait(smem_u32(&bar_acc_free[j & 1]), (((uint32_t)j >> 1) & 1u) ^ 1u);
        tc_fence_after();
#pragma unroll
        for (int sl = 0; sl < 2; ++sl)   // (d r0) W_0^T, columns of chunk j
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16(tmem + acc, umma_desc(s_base + kPbSmemG0 + sl * CN_SLAB_BYTES + k * kKmajKstep, kKmajLbo, kSboBytes),
                     umma_desc(a + kOffW0 + sl * 8192 + k * kKmajKstep, kKmajLbo, kSboBytes), idesc64, (sl | k) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 4; ++k)      // (d o1) W_sc^T
          umma_f16(tmem + acc + 64, umma_desc(s_base + kPbSmemGs + k * kKmajKstep, kKmajLbo, kSboBytes),
                   umma_desc(a + kOffSc + k * kKmajKstep, kKmajLbo, kSboBytes), idesc64, k ? 1u : 0u);
        umma_commit(smem_u32(&bar_acc_full[j & 1]));
        umma_commit(smem_u32(&bar_empty[st]));
      }
    }
  } else {
    const uint32_t quad = warp & 3u, row = quad * 32u + lane, lane_off = (quad * 32u) << 16;
    const bool leader = warp == 2 && lane == 0;
    const int64_t pose = ptile * 128 + row;
    const uint8_t *S = saved + ptile * (int64_t)kSavedTile + row * 128u;
    uint8_t *Gs = smem + kPbSmemGs;
    uint32_t pht = 0;
    auto saved_chunk = [&](uint32_t slab, uint32_t c8, float (&a)[8]) {
      const uint4 o = ldg_nc_u4(S + slab * CN_SLAB_BYTES + ((c8 ^ (row & 7u)) << 4));
      a[0] = bf16_lo(o.x); a[1] = bf16_hi(o.x); a[2] = bf16_lo(o.y); a[3] = bf16_hi(o.y);
      a[4] = bf16_lo(o.z); a[5] = bf16_hi(o.z); a[6] = bf16_lo(o.w); a[7] = bf16_hi(o.w);
    };
    auto put = [&](uint8_t *tile, uint32_t slab, uint32_t c8, const float (&a)[8]) {
      *reinterpret_cast<uint4 *>(tile + slab * CN_SLAB_BYTES + row * 128u + ((c8 ^ (row & 7u)) << 4)) =
          make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
    };
    auto gemm = [&](uint32_t d_col, uint32_t w_off, uint32_t idesc) {   // D = Gs [128 x 64] . W^T image
      fence_proxy_async_smem();
      tc_fence_before();
      named_bar_sync(1, 128);
      if (leader) {
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_f16(tmem + d_col, umma_desc(s_base + kPbSmemGs + k * kKmajKstep, kKmajLbo, kSboBytes),
                   umma_desc(s_base + kPbSmemW + w_off + k * kKmajKstep, kKmajLbo, kSboBytes), idesc, k ? 1u : 0u);
        umma_commit(smem_u32(&bar_tail));
      }
      mbar_wait(smem_u32(&bar_tail), pht & 1u);
      ++pht;
      tc_fence_after();
    };
    mbar_wait(smem_u32(&bar_w), 0);
    // B1: d o2 = gy w_o [o2 > 0]  (fp32 in registers: it is also the residual's gradient)
    const float gyv = pose < g.n_poses ? gy[pose] : 0.f;
    float d_o[64];
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) {
      float m[8], a[8];
      saved_chunk(4, c8, m);
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = d_o[c8 * 8 + i] = m[i] > 0.f ? gyv * s_wo[c8 * 8 + i] : 0.f;
      put(Gs, 0, c8, a);
    }
    gemm(0, img(128, 64) + img(64, 64), idesc64);          // (d o2) M_1^T
    // B2: d(pre-act of the block's hidden layer) = . act'(act(r2))
    {
      uint32_t v[64];
      tmem_ld32(tmem + lane_off, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tmem_ld32(tmem + lane_off + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tmem_wait_ld();
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8) {
        float m[8], a[8];
        saved_chunk(3, c8, m);
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = __uint_as_float(v[c8 * 8 + i]) * act_der(m[i], g.elu);
        put(Gs, 0, c8, a);
      }
    }
    gemm(64, img(128, 64), idesc64);                       // . M_0^T
    // B3: d o1 = d o2 + . act'(act(o1))
    {
      uint32_t v[64];
      tmem_ld32(tmem + lane_off + 64, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tmem_ld32(tmem + lane_off + 96, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tmem_wait_ld();
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8) {
        float m[8], a[8];
        saved_chunk(2, c8, m);
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = d_o[c8 * 8 + i] + __uint_as_float(v[c8 * 8 + i]) * act_der(m[i], g.elu);
        put(Gs, 0, c8, a);                                   // stays: A operand of the shortcut's chunks
      }
    }
    gemm(128, 0, idesc128);                                // (d o1) W_1^T  (128 columns)
    // B4: d r0 = . act'(act(r0)) -> operand tile G0
#pragma unroll 1
    for (int c4 = 0; c4 < 4; ++c4) {
      uint32_t v[32];
      tmem_ld32(tmem + lane_off + 128 + c4 * 32, v);
      tmem_wait_ld();
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float m[8], a[8];
        saved_chunk((uint32_t)c4 >> 1, (uint32_t)(c4 & 1) * 4 + q, m);
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = __uint_as_float(v[q * 8 + i]) * act_der(m[i], g.elu);
        put(smem + kPbSmemG0, (uint32_t)c4 >> 1, (uint32_t)(c4 & 1) * 4 + q, a);
      }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&bar_g));
    // ---- 42 chunks: d v_j = (d o1) W_sc^T + ((d r0) W_0^T) act'(act(v_j));  d(pre-act of c) = d v_j act'(v_j)
    for (int j = 0; j < kN5; ++j) {
      const uint32_t st = (uint32_t)j % kPbStages;
      const uint8_t *V = smem + st * kPbStage + row * 128u;
      mbar_wait(smem_u32(&bar_full[st]), ((uint32_t)j / kPbStages) & 1u);
      mbar_wait(smem_u32(&bar_acc_full[j & 1]), ((uint32_t)j >> 1) & 1u);
      tc_fence_after();
      uint32_t d0[64], d1[64];
      const uint32_t acc = tmem + lane_off + 256u + ((uint32_t)j & 1u) * 128u;
      tmem_ld32(acc, *reinterpret_cast<uint32_t(*)[32]>(&d0[0]));
      tmem_ld32(acc + 32, *reinterpret_cast<uint32_t(*)[32]>(&d0[32]));
      tmem_ld32(acc + 64, *reinterpret_cast<uint32_t(*)[32]>(&d1[0]));
      tmem_ld32(acc + 96, *reinterpret_cast<uint32_t(*)[32]>(&d1[32]));
      tmem_wait_ld();
      tc_fence_before();
      uint8_t *dst = dpc + (ptile * kN5 + j) * (int64_t)CN_SLAB_BYTES + row * 128u;
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8) {
        const uint32_t off = ((uint32_t)c8 ^ (row & 7u)) << 4;
        const uint4 vv = *reinterpret_cast<const uint4 *>(V + off), ee = *reinterpret_cast<const uint4 *>(V + CN_SLAB_BYTES + off);
        const float vf[8] = {bf16_lo(vv.x), bf16_hi(vv.x), bf16_lo(vv.y), bf16_hi(vv.y), bf16_lo(vv.z), bf16_hi(vv.z), bf16_lo(vv.w), bf16_hi(vv.w)};
        const float ef[8] = {bf16_lo(ee.x), bf16_hi(ee.x), bf16_lo(ee.y), bf16_hi(ee.y), bf16_lo(ee.z), bf16_hi(ee.z), bf16_lo(ee.w), bf16_hi(ee.w)};
        float a[8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
          a[i] = (__uint_as_float(d1[c8 * 8 + i]) + __uint_as_float(d0[c8 * 8 + i]) * act_der(ef[i], g.elu)) * act_der(vf[i], g.elu);
        *reinterpret_cast<uint4 *>(dst + off) =
            make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
      }
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(smem_u32(&bar_acc_free[j & 1]));
        mbar_arrive(smem_u32(&bar_empty[st]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------ point backward
constexpr uint32_t kPqSmemW = 0;                        // pack [kHbDs, kHbL1): W_s^T (4 x 16 KiB), W_c^T (32 KiB)
constexpr uint32_t kPqSmemAcat = kHbL1 - kHbDs;         // saved [a_1 .. a_4] tile, overwritten by d(pre-act) in place
constexpr uint32_t kPqSmemA0 = kPqSmemAcat + kAcatTile; // d(pre-act of c) operand tile [128 x 64]
constexpr uint32_t kPqSmemBytes = kPqSmemA0 + CN_SLAB_BYTES + 1024;
static_assert(kPqSmemAcat == 98304, "point backward weights");

__global__ void __launch_bounds__(kPtThreads, 1)
grasp_point_bwd_kernel(const uint8_t *__restrict__ dpc, const uint8_t *__restrict__ acat, const uint8_t *__restrict__ pack,
                       HeadGeom g, uint8_t *__restrict__ dtap, int64_t dtap_stride) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_w, bar_acat, bar_1, bar_2a, bar_2b;
  __shared__ uint32_t tmem_slot;
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t s_base = smem_u32(smem);
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
  if (tid == 0) {
    mbar_init(smem_u32(&bar_w), 1);
    mbar_init(smem_u32(&bar_acat), 1);
    mbar_init(smem_u32(&bar_1), 1);
    mbar_init(smem_u32(&bar_2a), 1);
    mbar_init(smem_u32(&bar_2b), 1);
    fence_barrier_init();
    mbar_expect_tx(smem_u32(&bar_w), kHbL1 - kHbDs);
    bulk_g2s(s_base + kPqSmemW, pack + kHbDs, kHbL1 - kHbDs, smem_u32(&bar_w));
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t quad = warp & 3u, half = warp >> 2, row = quad * 32u + lane, lane_off = (quad * 32u) << 16;
  constexpr uint32_t idesc256 = umma_idesc(128, 256, 0, 0), idesc128 = umma_idesc(128, 128, 0, 0);
  uint32_t ph = 0;
  bool first = true;
  for (int64_t tile = blockIdx.x; tile < g.n_tiles; tile += gridDim.x, ++ph) {
    if (tid == 0) {
      mbar_expect_tx(smem_u32(&bar_acat), kAcatTile);
      bulk_g2s(s_base + kPqSmemAcat, acat + tile * (int64_t)kAcatTile, kAcatTile, smem_u32(&bar_acat));
    }
    // d(pre-act of c) rows of this tile, gathered from the slabs of (pose tile, offset); padding rows are zero
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t q = tid + 256u * i, rr = q >> 3, ch = q & 7u;
      int pose, j;
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (row_pose(g, tile * 128 + rr, pose, j))
        v = ldg_nc_u4(dpc + ((int64_t)(pose >> 7) * kN5 + j) * CN_SLAB_BYTES + (uint32_t)(pose & 127) * 128u +
                      ((ch ^ ((uint32_t)pose & 7u)) << 4));
      *reinterpret_cast<uint4 *>(smem + kPqSmemA0 + rr * 128u + ((ch ^ (rr & 7u)) << 4)) = v;
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
      if (first) mbar_wait(smem_u32(&bar_w), 0);
      tc_fence_after();
#pragma unroll
      for (int k = 0; k < 4; ++k)        // d cat [128 x 256] = d(pre-act c) W_c^T
        umma_f16(tmem, umma_desc(s_base + kPqSmemA0 + k * kKmajKstep, kKmajLbo, kSboBytes),
                 umma_desc(s_base + kPqSmemW + (kHbComb - kHbDs) + k * kKmajKstep, kKmajLbo, kSboBytes), idesc256, k ? 1u : 0u);
      umma_commit(smem_u32(&bar_1));
    }
    first = false;
    mbar_wait(smem_u32(&bar_1), ph & 1u);
    mbar_wait(smem_u32(&bar_acat), ph & 1u);
    tc_fence_after();
    // d(pre-act of a_s) = d cat_s act'(a_s), in place over a_s
    for (int ss = 0; ss < 2; ++ss) {
      const int s = 2 * (int)half + ss;
      uint32_t v[64];
      tmem_ld32(tmem + lane_off + s * 64, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tmem_ld32(tmem + lane_off + s * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tmem_wait_ld();
      uint8_t *p = smem + kPqSmemAcat + s * CN_SLAB_BYTES + row * 128u;
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8) {
        uint4 *q = reinterpret_cast<uint4 *>(p + (((uint32_t)c8 ^ (row & 7u)) << 4));
        const uint4 o = *q;
        const float m[8] = {bf16_lo(o.x), bf16_hi(o.x), bf16_lo(o.y), bf16_hi(o.y), bf16_lo(o.z), bf16_hi(o.z), bf16_lo(o.w), bf16_hi(o.w)};
        float a[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = __uint_as_float(v[c8 * 8 + i]) * act_der(m[i], g.elu);
        *q = make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
      }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      for (int s = 0; s < 4; ++s) {      // d tap_s [128 x 128] = d(pre-act a_s) W_s^T;  taps 0, 1 beside d cat, taps 2, 3 over it
        const uint32_t d = s < 2 ? 256u + (uint32_t)s * 128u : (uint32_t)(s - 2) * 128u;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_f16(tmem + d, umma_desc(s_base + kPqSmemAcat + s * CN_SLAB_BYTES + k * kKmajKstep, kKmajLbo, kSboBytes),
                   umma_desc(s_base + kPqSmemW + s * img(128, 64) + k * kKmajKstep, kKmajLbo, kSboBytes), idesc128, k ? 1u : 0u);
        if (s == 1) umma_commit(smem_u32(&bar_2a));
      }
      umma_commit(smem_u32(&bar_2b));
    }
    for (int s = 0; s < 4; ++s) {
      if (s == 0) mbar_wait(smem_u32(&bar_2a), ph & 1u);
      if (s == 2) mbar_wait(smem_u32(&bar_2b), ph & 1u);
      tc_fence_after();
      const uint32_t d = s < 2 ? 256u + (uint32_t)s * 128u : (uint32_t)(s - 2) * 128u;
      uint32_t v[64];
      tmem_ld32(tmem + lane_off + d + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tmem_ld32(tmem + lane_off + d + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tmem_wait_ld();
      uint8_t *dst = dtap + s * dtap_stride + tile * (int64_t)CN_ACT_TILE_BYTES + half * CN_SLAB_BYTES + row * 128u;
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8)
        *reinterpret_cast<uint4 *>(dst + (((uint32_t)c8 ^ (row & 7u)) << 4)) =
            make_uint4(cvt_bf16x2(__uint_as_float(v[c8 * 8 + 0]), __uint_as_float(v[c8 * 8 + 1])),
                       cvt_bf16x2(__uint_as_float(v[c8 * 8 + 2]), __uint_as_float(v[c8 * 8 + 3])),
                       cvt_bf16x2(__uint_as_float(v[c8 * 8 + 4]), __uint_as_float(v[c8 * 8 + 5])),
                       cvt_bf16x2(__uint_as_float(v[c8 * 8 + 6]), __uint_as_float(v[c8 * 8 + 7])));
    }
    tc_fence_before();
    __syncthreads();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// g (bf16 tile images) += src (bf16 tile images): one rounding per element
__global__ void __launch_bounds__(256) tile_add_tiles_kernel(uint4 *__restrict__ gdst, const uint4 *__restrict__ src, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const uint4 a = gdst[i], b = ldg_nc_u4(src + i);
  auto add = [](uint32_t x, uint32_t y) { return pack_bf16x2(bf16_lo(x) + bf16_lo(y), bf16_hi(x) + bf16_hi(y)); };
  gdst[i] = make_uint4(add(a.x, b.x), add(a.y, b.y), add(a.z, b.z), add(a.w, b.w));
}

static int head_geom(HeadGeom &g, int64_t n_tiles, int b, int n_pad, int np, int activation, const char *who) {
  CN_REQUIRE(b > 0 && np > 0 && n_pad > 0 && n_pad % 128 == 0 && np * kN5 <= n_pad, "%s: b=%d n_pad=%d np=%d", who, b, n_pad, np);
  CN_REQUIRE(n_tiles == (int64_t)b * (n_pad / 128), "%s: n_tiles %lld != b * n_pad / 128", who, (long long)n_tiles);
  CN_REQUIRE(activation == 0 || activation == 1, "%s: activation %d (0 relu, 1 elu)", who, activation);
  g.n_tiles = n_tiles;
  g.n_pad = n_pad;
  g.np = np;
  g.n_poses = b * np;
  g.elu = activation;
  return 0;
}

template <typename K> static int set_smem(K kernel, uint32_t bytes, const char *who) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) {
    set_error("%s: cudaFuncSetAttribute: %s", who, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

}  // namespace cn

extern "C" int cn_grasp_head_pack(const float *params, void *pack, void *stream) {
  using namespace cn;
  CN_REQUIRE(params && pack, "cn_grasp_head_pack: null pointer");
  CN_REQUIRE(((uintptr_t)pack & 127) == 0, "cn_grasp_head_pack: pack must be 128-byte aligned");
  PackTable t{};
  int n = 0;
  // forward image B[n = out][k = in] reads kernel[k][n]; backward image B[n = in][k = out] reads kernel[n][k]
  auto both = [&](int src, int in, int out, uint32_t fwd, uint32_t bwd) {
    t.e[n++] = PackEntry{src, 1, out, out, in, fwd};
    t.e[n++] = PackEntry{src, out, 1, in, out, bwd};
  };
  for (int s = 0; s < 4; ++s) both(kHpDs + s * kHpDsStride, 128, 64, kHfDs + s * img(64, 128), kHbDs + s * img(128, 64));
  both(kHpComb, 256, 64, kHfComb, kHbComb);
  both(kHpL0, 2688, 128, kHfL0, kHbL0);
  both(kHpL1, 128, 64, kHfL1, kHbL1);
  both(kHpSc, 2688, 64, kHfSc, kHbSc);
  both(kHpM0, 64, 64, kHfM0, kHbM0);
  both(kHpM1, 64, 64, kHfM1, kHbM1);
  t.n = n;
  grasp_head_pack_kernel<<<dim3(32, n + 1), 256, 0, (cudaStream_t)stream>>>(params, (uint8_t *)pack, t);
  return check_launch("cn_grasp_head_pack");
}

extern "C" int cn_grasp_head_fwd(const void *tap_tiles, int64_t n_tiles, int32_t b, int32_t n_pad, int32_t np, const void *pack,
                                 int32_t activation, void *vslabs, void *acat, void *saved, float *y, void *stream) {
  using namespace cn;
  CN_REQUIRE(tap_tiles && pack && vslabs && saved && y, "cn_grasp_head_fwd: null pointer");
  CN_REQUIRE((((uintptr_t)tap_tiles | (uintptr_t)pack | (uintptr_t)vslabs | (uintptr_t)acat | (uintptr_t)saved) & 127) == 0,
             "cn_grasp_head_fwd: buffers must be 128-byte aligned");
  HeadGeom g;
  if (int rc = head_geom(g, n_tiles, b, n_pad, np, activation, "cn_grasp_head_fwd")) return rc;
  if (int rc = set_smem(grasp_point_fwd_kernel, kPfSmemBytes, "cn_grasp_head_fwd")) return rc;
  if (int rc = set_smem(grasp_pose_fwd_kernel, kPsSmemBytes, "cn_grasp_head_fwd")) return rc;
  const unsigned grid = (unsigned)(n_tiles < num_sms() ? n_tiles : num_sms());
  grasp_point_fwd_kernel<<<grid, kPtThreads, kPfSmemBytes, (cudaStream_t)stream>>>(
      (const uint8_t *)tap_tiles, n_tiles * (int64_t)CN_ACT_TILE_BYTES, (const uint8_t *)pack, g, (uint8_t *)vslabs,
      (uint8_t *)acat);
  if (int rc = check_launch("cn_grasp_head_fwd (points)")) return rc;
  const unsigned p_tiles = (unsigned)((g.n_poses + 127) / 128);
  grasp_pose_fwd_kernel<<<p_tiles, kPoseThreads, kPsSmemBytes, (cudaStream_t)stream>>>(
      (const uint8_t *)vslabs, (const uint8_t *)pack, g, (uint8_t *)saved, y);
  return check_launch("cn_grasp_head_fwd (poses)");
}

extern "C" int cn_grasp_head_bwd(const float *gy, int64_t n_tiles, int32_t b, int32_t n_pad, int32_t np, const void *pack,
                                 int32_t activation, const void *vslabs, const void *acat, const void *saved, void *dpc_ws,
                                 void *dtap_tiles, void *stream) {
  using namespace cn;
  CN_REQUIRE(gy && pack && vslabs && acat && saved && dpc_ws && dtap_tiles, "cn_grasp_head_bwd: null pointer");
  CN_REQUIRE((((uintptr_t)pack | (uintptr_t)vslabs | (uintptr_t)acat | (uintptr_t)saved | (uintptr_t)dpc_ws |
               (uintptr_t)dtap_tiles) & 127) == 0, "cn_grasp_head_bwd: buffers must be 128-byte aligned");
  HeadGeom g;
  if (int rc = head_geom(g, n_tiles, b, n_pad, np, activation, "cn_grasp_head_bwd")) return rc;
  if (int rc = set_smem(grasp_pose_bwd_kernel, kPbSmemBytes, "cn_grasp_head_bwd")) return rc;
  if (int rc = set_smem(grasp_point_bwd_kernel, kPqSmemBytes, "cn_grasp_head_bwd")) return rc;
  const unsigned p_tiles = (unsigned)((g.n_poses + 127) / 128);
  grasp_pose_bwd_kernel<<<p_tiles, kPoseThreads, kPbSmemBytes, (cudaStream_t)stream>>>(
      gy, (const uint8_t *)vslabs, (const uint8_t *)saved, (const uint8_t *)pack, g, (uint8_t *)dpc_ws);
  if (int rc = check_launch("cn_grasp_head_bwd (poses)")) return rc;
  const unsigned grid = (unsigned)(n_tiles < num_sms() ? n_tiles : num_sms());
  grasp_point_bwd_kernel<<<grid, kPtThreads, kPqSmemBytes, (cudaStream_t)stream>>>(
      (const uint8_t *)dpc_ws, (const uint8_t *)acat, (const uint8_t *)pack, g, (uint8_t *)dtap_tiles,
      n_tiles * (int64_t)CN_ACT_TILE_BYTES);
  return check_launch("cn_grasp_head_bwd (points)");
}

// ---- cn_grasp_head_bwd without a caller-provided scratch buffer (no counterpart in the reference; added
// for callers that cannot allocate: code under CUDA-graph capture, serving loops).
// The d(pre-activation of c) slabs live in a file-scope device array instead: CN_HEAD_BWD_STATIC_MAX_POSES poses =
// 8 pose tiles x 42 x 16 KiB = 5.25 MiB, held by every GPU that loads the library until the process exits.
namespace cn {
constexpr int64_t kStaticDpcBytes = (int64_t)(CN_HEAD_BWD_STATIC_MAX_POSES / 128) * kN5 * CN_SLAB_BYTES;
static __device__ __align__(256) uint8_t g_head_bwd_dpc[kStaticDpcBytes];

// In host code `g_head_bwd_dpc` is only a stand-in for the symbol, NOT a device address: the address is asked for
// once per device, with that device current, and cached.
constexpr int kMaxDevices = 64;
static std::mutex g_dpc_mutex;
static void *g_dpc_ptr[kMaxDevices] = {};

static int static_dpc_address(void **out, const char *who) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    set_error("%s: cudaGetDevice: %s (device %d)", who, cudaGetErrorString(e), dev);
    return e != cudaSuccess ? (int)e : -1;
  }
  std::lock_guard<std::mutex> lock(g_dpc_mutex);
  if (!g_dpc_ptr[dev]) {
    void *ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_head_bwd_dpc);
    if (e != cudaSuccess || !ptr) {
      set_error("%s: cudaGetSymbolAddress(g_head_bwd_dpc) on device %d: %s", who, dev, cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : -1;
    }
    g_dpc_ptr[dev] = ptr;
  }
  *out = g_dpc_ptr[dev];
  return 0;
}
}  // namespace cn

extern "C" int cn_grasp_head_bwd_static(const float *gy, int64_t n_tiles, int32_t b, int32_t n_pad, int32_t np,
                                        const void *pack, int32_t activation, const void *vslabs, const void *acat,
                                        const void *saved, void *dtap_tiles, void *stream) {
  using namespace cn;
  CN_REQUIRE(b > 0 && np > 0, "cn_grasp_head_bwd_static: b=%d np=%d", b, np);
  CN_REQUIRE((int64_t)b * np <= CN_HEAD_BWD_STATIC_MAX_POSES,
             "cn_grasp_head_bwd_static: %lld poses exceed the static scratch (CN_HEAD_BWD_STATIC_MAX_POSES = %d); "
             "use cn_grasp_head_bwd with a caller-provided dpc_ws", (long long)b * np, CN_HEAD_BWD_STATIC_MAX_POSES);
  void *dpc = nullptr;
  if (int rc = static_dpc_address(&dpc, "cn_grasp_head_bwd_static")) return rc;
  // every slab row the point kernel reads is written by the pose kernel of the SAME call (rows of valid poses only):
  // nothing in the array has to start at zero, the previous call's contents are simply overwritten
  return cn_grasp_head_bwd(gy, n_tiles, b, n_pad, np, pack, activation, vslabs, acat, saved, dpc, dtap_tiles, stream);
}

extern "C" int cn_tile_add_tiles(void *g_tiles, const void *src_tiles, int64_t n_tiles, void *stream) {
  using namespace cn;
  CN_REQUIRE(g_tiles && src_tiles, "cn_tile_add_tiles: null pointer");
  CN_REQUIRE(n_tiles >= 0, "cn_tile_add_tiles: n_tiles %lld", (long long)n_tiles);
  CN_REQUIRE((((uintptr_t)g_tiles | (uintptr_t)src_tiles) & 15) == 0, "cn_tile_add_tiles: buffers must be 16-byte aligned");
  if (n_tiles == 0) return 0;
  const int64_t n = n_tiles * (CN_ACT_TILE_BYTES / 16);
  tile_add_tiles_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>((uint4 *)g_tiles, (const uint4 *)src_tiles, n);
  return check_launch("cn_tile_add_tiles");
}
