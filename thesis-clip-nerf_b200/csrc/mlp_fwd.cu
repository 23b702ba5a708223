// Fused MLP forward on tcgen05 / TMEM (MVResNetMLPNeRFEmbedding + RenderReadout, layers.py:334-397).
//
// Two launches of one warp-specialised, persistent kernel template:
//   kFusion = false : per-view stage -- Dense 379(384)->128 and 3 pre-activation ResNet blocks
//                     on the MLP-input tiles written by cn_project_gather_fwd
//   kFusion = true  : mean over views, 3 ResNet blocks, readout 128->4, sigmoid / softplus
//
// One CTA per SM works on PAIRS of 128-row tiles so that the tensor pipe works on one tile while
// the epilogue warps work on the other (ping-pong):
//   warp 0      TMA producer   : streams X slabs and weight slabs (16 KiB each) through an
//                                8-slot shared-memory ring with cp.async.bulk + mbarriers
//   warp 1      MMA issuer     : one thread issues tcgen05.mma (M=128, N=128, K=16, bf16 -> fp32)
//   warps 2-9   epilogue tile 0: TMEM -> registers -> bias / ReLU -> bf16 -> swizzled smem operand
//   warps 10-17 epilogue tile 1  (8 warps per tile: 4 TMEM lane quadrants x 2 column halves)
// The fusion stage stages the nv per-view tiles of each 128-row tile in shared memory with TMA and
// averages them from there.
// TMEM (512 columns): per tile a 128-column fp32 residual stream `h` and a 128-column `tmp`.
// The second GEMM of every block accumulates straight into `h`, so the residual add is free and
// the stream stays fp32; Dense biases of the residual branch are applied as cumulative bias
// vectors when `h` is read (tables built by cn_pack_weights).
// Scheduling (tools/trace_mlp_chain.py): a layer of one tile is a serial chain -- 8 MMAs issued
// (~600 cycles), epilogue wake-up (~250), conversion (~750), publish + MMA wake-up (~350) -- so the pair advances one layer
// per ~2 000 cycles.  Three things shorten or hide parts of it: the two accumulators of a tile swap roles every pair
// (the first Dense of the next pair runs under this pair's readout), the first Dense staggers the two tiles by two K
// blocks (their epilogues alternate), and in the fusion stage the operand tile is handed over per K slab.
#include "common.cuh"
#include "mlp_layout.cuh"

namespace cn {

constexpr int kFwdThreads = 608;                                   // 2 + 16 warps + the X producer warp
constexpr int kEpiThreads = 256;                                   // epilogue threads per tile
constexpr int kEpiWarps = kEpiThreads / 32;
constexpr int kMaxStageViews = 3;
// First Dense of a pair: order of its 12 (K block, tile) steps.  Tile 0 runs two K blocks ahead of tile 1, so tile 1
// enters the blocks ~half a layer period after tile 0: the two tiles' epilogues then ALTERNATE (each overlaps the other
// tile's GEMM) instead of running at the same time on the same schedulers.  An offset between the tiles of more than one
// GEMM is self-sustaining under in-order issue (the issuer waits for the later tile), so it only has to be created here.
// The X producer loads in the same order; a weight slab is released after tile 1 has used it (3 of 4 ring slots live).
__device__ constexpr int8_t kL0StepKb[12] = {0, 1, 2, 0, 3, 1, 4, 2, 5, 3, 4, 5};
__device__ constexpr int8_t kL0StepT[12] = {0, 0, 0, 1, 0, 1, 0, 1, 0, 1, 1, 1};                                  // views staged in smem by the fusion prologue
constexpr uint32_t kSmemAct = 0;                                   // 2 x 32 KiB
constexpr uint32_t kSmemRing = 2 * CN_ACT_TILE_BYTES;
constexpr uint32_t kSmemTabBytes = 7 * 128 * 4 + 512 * 4 + 16;
constexpr uint32_t kSmemTabBytesChain = 14 * 128 * 4 + 512 * 4 + 16;   // both stages' tables (chained kernel)
// Rings of 16 KiB slabs.  Fusion stage: ONE 8-slot ring (per-view tiles + weights).  Per-view stage: a 4-slot weight
// ring and a 5-slot ring of X slabs with its own producer warp, so that the first X slabs of the NEXT pair stream in
// from HBM while the blocks of this pair run.  tools/trace_mlp_fwd.py: the first Dense of a pair still takes ~9000 of
// the pair's ~28000 cycles -- 80 KiB in flight against ~3500 cycles of loaded HBM latency is 23 B/clk, the SM's fair
// share of HBM; the stage moves 224 KiB per tile through HBM (X in, 3 activation tiles + h3 out) and is bound by that.
template <bool kFusion, bool kChain = false> struct FwdSmem {
  static constexpr int kRingSlots = kFusion ? 8 : 4;
  static constexpr int kXSlots = kFusion ? 0 : 5;
  static constexpr uint32_t kXRing = kSmemRing + kRingSlots * CN_SLAB_BYTES;
  static constexpr uint32_t kTab = kXRing + kXSlots * CN_SLAB_BYTES;
  static constexpr uint32_t kBytes = kTab + (kChain ? kSmemTabBytesChain : kSmemTabBytes) + 1024;  // + alignment slack
};

struct FwdBars {
  uint64_t full[8];
  uint64_t empty[8];
  uint64_t xfull[8];
  uint64_t xempty[8];
  uint64_t h_ready[2];      // second Dense of a block committed (residual stream updated)
  uint64_t h0_ready[2];     // first Dense of the pair committed.  Its own barrier: it is issued right behind the last block GEMM
                            // of the previous pair (no epilogue step in between), and two completions of ONE barrier that
                            // the waiter has not observed in between would alias its phase parity
  uint64_t tmp_ready[2];
  uint64_t act_ready[2];    // K slab 1 of ACT[t] written (the whole tile is then in shared memory)
  uint64_t act_ready0[2];   // K slab 0 of ACT[t] written: the next Dense starts on it while the epilogue converts slab 1
  uint64_t views_ready[2];  // fusion stage: the ring slots holding tile t's per-view slabs have landed
};

// TMEM accumulator (this thread's lane) + bias -> (ReLU) -> bf16 -> swizzled activation tile in shared memory.  The two
// warps of a lane quadrant split the columns so that each covers BOTH K slabs of the tile: warp `half` takes columns
// [32 half, 32 half + 32) of slab 0 first, then [64 + 32 half, ...) of slab 1.  When slab 0 is written the warp arrives on
// `bar_s0` (if given): the MMA thread issues the four MMAs of K slab 0 of the next Dense while the warps convert slab 1
// (tools/trace_mlp_chain.py: issuing a GEMM takes ~600 cycles, as long as the epilogue's second half).
// Both 32-column loads are issued before the wait.
// `tap` (optional): this row's 128 fp32 pre-activation values (h + bias) are also written there -- the stage
// outputs MVResNetMLPNeRFEmbedding(complete_output=True) returns (layers.py:364-377).
// `tap16` (optional): the same values as a bf16 activation-tile image (this thread writes its row's 128-byte line).
// `mask` (optional, kRelu): tile whose sign pattern replaces the ReLU (Jacobian-vector product with the primal masks).
// kPhased: the 32 columns of slab 1 are loaded from TMEM only after slab 0 has been converted and signalled -- half the
// live registers (the kernels sit at their 96-register cap), the second load's latency hides under the MMAs just released.
template <bool kRelu, bool kPhased = false>
__device__ __forceinline__ void epi_tmem_to_act(uint32_t taddr, const float *__restrict__ bias, uint8_t *act,
                                                uint32_t row, uint32_t half, uint32_t bar_s0,
                                                float *__restrict__ tap = nullptr, uint8_t *__restrict__ tap16 = nullptr,
                                                const uint8_t *__restrict__ mask = nullptr) {
  uint32_t r[64];
  tmem_ld32(taddr + half * 32, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
  if (!kPhased) tmem_ld32(taddr + 64 + half * 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
  tmem_wait_ld();
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    if (kPhased && q == 4) {
      tmem_ld32(taddr + 64 + half * 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
      tmem_wait_ld();
    }
    const uint32_t col = (uint32_t)(q >> 2) * 64 + half * 32 + (uint32_t)(q & 3) * 8;   // first column of this 8-column chunk
    const float4 b0 = *reinterpret_cast<const float4 *>(bias + col);
    const float4 b1 = *reinterpret_cast<const float4 *>(bias + col + 4);
    float v[8];
    v[0] = __uint_as_float(r[q * 8 + 0]) + b0.x; v[1] = __uint_as_float(r[q * 8 + 1]) + b0.y;
    v[2] = __uint_as_float(r[q * 8 + 2]) + b0.z; v[3] = __uint_as_float(r[q * 8 + 3]) + b0.w;
    v[4] = __uint_as_float(r[q * 8 + 4]) + b1.x; v[5] = __uint_as_float(r[q * 8 + 5]) + b1.y;
    v[6] = __uint_as_float(r[q * 8 + 6]) + b1.z; v[7] = __uint_as_float(r[q * 8 + 7]) + b1.w;
    if (tap) {
      *reinterpret_cast<float4 *>(tap + col) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4 *>(tap + col + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    if (tap16)
      *reinterpret_cast<uint4 *>(tap16 + tile_chunk_off(row, col >> 3)) =
          make_uint4(cvt_bf16x2(v[0], v[1]), cvt_bf16x2(v[2], v[3]), cvt_bf16x2(v[4], v[5]), cvt_bf16x2(v[6], v[7]));
    uint4 o;
    if (kRelu && mask) {
      // tangent pass: the activation's derivative is the PRIMAL mask, read from the tile the forward saved
      const uint4 m = ldg_nc_u4(mask + tile_chunk_off(row, col >> 3));
      o = make_uint4(mul_bf16x2(cvt_bf16x2(v[0], v[1]), gt0_bf16x2(m.x)), mul_bf16x2(cvt_bf16x2(v[2], v[3]), gt0_bf16x2(m.y)),
                     mul_bf16x2(cvt_bf16x2(v[4], v[5]), gt0_bf16x2(m.z)), mul_bf16x2(cvt_bf16x2(v[6], v[7]), gt0_bf16x2(m.w)));
    } else if (kRelu)  // relu folded into the conversion (F2FP.RELU)
      o = make_uint4(cvt_relu_bf16x2(v[0], v[1]), cvt_relu_bf16x2(v[2], v[3]), cvt_relu_bf16x2(v[4], v[5]),
                     cvt_relu_bf16x2(v[6], v[7]));
    else
      o = make_uint4(cvt_bf16x2(v[0], v[1]), cvt_bf16x2(v[2], v[3]), cvt_bf16x2(v[4], v[5]), cvt_bf16x2(v[6], v[7]));
    *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, col >> 3)) = o;
    if (q == 3 && bar_s0) {                       // K slab 0 of this warp's rows is in shared memory
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(bar_s0);
    }
  }
}

// kExt: the pose-query extensions (bf16 tap tiles, saved hidden activations, tangent pass with external masks) are
// compiled into their own instantiations -- the render path's kernels carry none of their predicates or registers.
// kChain (per-view stage only, nv = 1): with ONE source view the mean over views is the identity, so the fusion stage
// is chained onto the per-view stage inside the same kernel -- six blocks and the readout per tile; the bf16 per-view
// output never leaves the SM (it is rounded to bf16 in registers exactly as the two-kernel path stores it, so the
// results are bit-identical), the fusion stage's lock-step prologue disappears, and both stages' saved activation
// tiles are written in the layouts the backward kernels already read.
template <bool kFusion, bool kExt, bool kChain = false>
__global__ void __launch_bounds__(kFwdThreads, 1)
mlp_fwd_kernel(const uint8_t *__restrict__ packed, const uint8_t *__restrict__ in_tiles, int64_t n_tiles,
               int nv, int64_t tiles_per_view, uint8_t *__restrict__ out_tiles, float *__restrict__ rgbs,
               uint8_t *__restrict__ acts, float *__restrict__ taps, int tap_bf16_arg, int save_hidden_arg,
               const uint8_t *__restrict__ mask_acts_arg, uint8_t *__restrict__ acts_u = nullptr,
               const uint8_t *__restrict__ mask_acts_u_arg = nullptr) {
  static_assert(!(kFusion && kChain), "the chained kernel is the per-view kernel continued");
  using L = FwdSmem<kFusion, kChain>;
  const int tap_bf16 = kExt ? tap_bf16_arg : 0, save_hidden = kExt ? save_hidden_arg : 0;
  const uint8_t *const mask_acts = kExt ? mask_acts_arg : nullptr;
  const uint8_t *const mask_acts_u = kExt ? mask_acts_u_arg : nullptr;
  constexpr int kNumBlk = kChain ? 6 : 3;                 // blocks per tile
  constexpr int kRoOff = (kChain ? 14 : 7) * 128;         // readout weights behind the bias tables
  // Epilogue -> MMA hand-off per K slab (epi_tmem_to_act), with the TMEM loads phased: fusion stage only.  Measured on the
  // per-view and chained kernels too (tools/bench_mlp_fwd.py): with phased loads they no longer spill more, but the
  // chained kernel without saves still ran 2.5 % slower and the per-view stage 4 % slower (their tiles are already
  // staggered by the first Dense), the fusion stage 5 ... 9 % faster.
  constexpr bool kSlabSplit = kFusion;
  constexpr bool kPhased = kFusion;
  constexpr int kRingSlots = L::kRingSlots;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) FwdBars bars;
  __shared__ uint32_t tmem_slot;

  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t s_base = smem_u32(smem);
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  float *tab = reinterpret_cast<float *>(smem + L::kTab);
  constexpr int kActsPerTile = kFusion ? 4 : 3;  // relu(h) of every block (+ relu(e)); relu(t) is recomputed by the backward
  // pose-query training (double backward): the hidden activations relu(t + ba) are saved too (3 more tiles per tile),
  // and the TANGENT pass (mask_acts != nullptr) re-runs this kernel on tangent inputs with the saved tiles as masks
  // and without biases -- the Jacobian-vector product of the embedding with the primal ReLU pattern
  const int acts_stride = kActsPerTile + ((save_hidden || mask_acts) ? 3 : 0);
  constexpr int kBlk0 = kFusion ? 3 : 0;
  const bool staged = kFusion && nv <= kMaxStageViews;  // view tiles go through the ring (else: global loads)
  const uint32_t view_slots = staged ? 2u * (uint32_t)nv : 0u;

  // ---- one-time setup
  {
    const float *gt = reinterpret_cast<const float *>(packed + kPackTables);
    for (int i = threadIdx.x; i < 7 * 128; i += kFwdThreads)
      tab[i] = mask_acts ? 0.f : gt[(kFusion ? kTabFusion : kTabFeature) + i];
    if (kChain)
      for (int i = threadIdx.x; i < 7 * 128; i += kFwdThreads) tab[7 * 128 + i] = mask_acts ? 0.f : gt[kTabFusion + i];
    if (kFusion || kChain)
      for (int i = threadIdx.x; i < 516; i += kFwdThreads) tab[kRoOff + i] = gt[kTabReadoutW + i];
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < kRingSlots; ++i) {
      mbar_init(smem_u32(&bars.full[i]), 1);
      mbar_init(smem_u32(&bars.empty[i]), 1);
    }
    for (int i = 0; i < L::kXSlots; ++i) {
      mbar_init(smem_u32(&bars.xfull[i]), 1);
      mbar_init(smem_u32(&bars.xempty[i]), 1);
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(smem_u32(&bars.h_ready[t]), 1);
      mbar_init(smem_u32(&bars.h0_ready[t]), 1);
      mbar_init(smem_u32(&bars.tmp_ready[t]), 1);
      mbar_init(smem_u32(&bars.act_ready[t]), kEpiWarps);   // one arrival per epilogue warp (behind a __syncwarp)
      mbar_init(smem_u32(&bars.act_ready0[t]), kEpiWarps);
      mbar_init(smem_u32(&bars.views_ready[t]), 1);
    }
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

  const int64_t n_pairs = (n_tiles + 1) / 2;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      uint32_t pc = 0;
      auto load = [&](const uint8_t *src) {
        const uint32_t slot = pc % kRingSlots;
        mbar_wait(smem_u32(&bars.empty[slot]), ((pc / kRingSlots) & 1u) ^ 1u);
        mbar_expect_tx(smem_u32(&bars.full[slot]), CN_SLAB_BYTES);
        bulk_g2s(s_base + kSmemRing + slot * CN_SLAB_BYTES, src, CN_SLAB_BYTES, smem_u32(&bars.full[slot]));
        ++pc;
      };
      auto load_weights = [&](int blk, int j) {
        for (int s = 0; s < 2; ++s) load(packed + pack_fwd_block(kBlk0 + blk, j) + s * CN_SLAB_BYTES);
      };
      // fusion stage: the nv per-view tiles of fusion tile `tile` go through the ring as 2 slabs each
      auto stage_views = [&](int64_t tile) {
        const int64_t b = tile / tiles_per_view, tt = tile % tiles_per_view;
        for (int v = 0; v < nv; ++v)
          for (int sl = 0; sl < 2; ++sl)
            load(in_tiles + ((b * nv + v) * tiles_per_view + tt) * (int64_t)CN_ACT_TILE_BYTES + sl * CN_SLAB_BYTES);
      };
      int64_t pi = 0;
      for (int64_t p = blockIdx.x; p < n_pairs; p += gridDim.x, ++pi) {
        const bool valid1 = 2 * p + 1 < n_tiles;
        if (!kFusion) {
          for (int kb = 0; kb < 6; ++kb) load(packed + kPackFwdW0 + kb * CN_SLAB_BYTES);
          CN_TRACE_AT(pi, 21);
          for (int blk = 0; blk < kNumBlk; ++blk)
            for (int j = 0; j < 2; ++j) load_weights(blk, j);
          CN_TRACE_AT(pi, 22);
        } else {
          if (staged) stage_views(2 * p);
          load_weights(0, 0);  // first weights before the second tile's views: MMA of tile 0 can start
          if (staged && valid1) stage_views(2 * p + 1);
          load_weights(0, 1);
          for (int blk = 1; blk < 3; ++blk)
            for (int j = 0; j < 2; ++j) load_weights(blk, j);
        }
      }
    }
  } else if (warp == 18) {
    // =========================== X producer (per-view stage) ===========================
    if (!kFusion && lane == 0) {
      constexpr int kXS = L::kXSlots > 0 ? L::kXSlots : 1;
      uint32_t xc = 0;
      int64_t pi = 0;
      for (int64_t p = blockIdx.x; p < n_pairs; p += gridDim.x, ++pi) {
        const int nt = (2 * p + 1 < n_tiles) ? 2 : 1;
#pragma unroll
        for (int st = 0; st < 12; ++st) {
            const int kb = kL0StepKb[st], t = kL0StepT[st];
            if (t >= nt) continue;
            const uint32_t slot = xc % kXS;
            mbar_wait(smem_u32(&bars.xempty[slot]), ((xc / kXS) & 1u) ^ 1u);
            mbar_expect_tx(smem_u32(&bars.xfull[slot]), CN_SLAB_BYTES);
            bulk_g2s(s_base + L::kXRing + slot * CN_SLAB_BYTES,
                     in_tiles + (2 * p + t) * (int64_t)CN_X_TILE_BYTES + kb * CN_SLAB_BYTES, CN_SLAB_BYTES,
                     smem_u32(&bars.xfull[slot]));
            ++xc;
            if (kb == 0 && t == 0) CN_TRACE_AT(pi, 20);
          }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(128, 128, 0, 0);
      uint32_t cc = 0;
      uint32_t n_act[2] = {0, 0};
      auto next_slot = [&]() -> uint32_t {
        const uint32_t slot = cc % kRingSlots;
        mbar_wait(smem_u32(&bars.full[slot]), (cc / kRingSlots) & 1u);
        ++cc;
        return slot;
      };
      auto slot_addr = [&](uint32_t slot) { return s_base + kSmemRing + slot * CN_SLAB_BYTES; };
      constexpr int kXS = L::kXSlots > 0 ? L::kXSlots : 1;
      uint32_t xcc = 0;
      auto next_x_slot = [&]() -> uint32_t {
        const uint32_t slot = xcc % kXS;
        mbar_wait(smem_u32(&bars.xfull[slot]), (xcc / kXS) & 1u);
        ++xcc;
        return slot;
      };
      // D(tmem) (+)= ACT[t] (2 slabs) . W (2 ring slots)
      auto gemm_act = [&](int t, uint32_t d_tmem, uint32_t s0, uint32_t s1, bool accumulate, uint32_t act_parity) {
        const uint32_t a_base = s_base + kSmemAct + t * CN_ACT_TILE_BYTES;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const uint64_t da = umma_desc(a_base + s * CN_SLAB_BYTES, kKmajLbo, kSboBytes);
          const uint64_t db = umma_desc(slot_addr(s ? s1 : s0), kKmajLbo, kSboBytes);
          // K slab s of the operand.  The slab-wise hand-off is compiled into the fusion stage only (-5 % there); the
          // chained kernel, at its 96-register cap, spilled twice as much with it and ran 9 % slower.
          if (kSlabSplit || s == 0) {
            mbar_wait(smem_u32((kSlabSplit && s == 0) ? &bars.act_ready0[t] : &bars.act_ready[t]), act_parity);
            tc_fence_after();
          }
#pragma unroll
          for (int j = 0; j < 4; ++j)
            umma_f16(d_tmem, da + j * kKmajKstepEnc, db + j * kKmajKstepEnc, idesc, (accumulate || s || j) ? 1u : 0u);
        }
      };
      int64_t pi = 0;
      for (int64_t p = blockIdx.x; p < n_pairs; p += gridDim.x, ++pi) {
        const int nt = (2 * p + 1 < n_tiles) ? 2 : 1;
        CN_TRACE_AT(pi, 0);
        // per-view / chained kernel: the two 128-column accumulators of a tile SWAP roles every pair.  The first Dense
        // of pair p+1 accumulates into the columns pair p used as `tmp` (idle since the last E_b, which the act_ready
        // wait of the last GEMM already covers) while the final epilogue of pair p is still reading the old `h`: the
        // X-bound first Dense overlaps the readout instead of waiting for it.  The old `h` columns become `tmp`, first
        // written by the block GEMM behind act_ready of E_a -- which the same epilogue warps signal after their readout.
        const uint32_t col_h = (!kFusion && (pi & 1)) ? 128u : 0u, col_tmp = 128u - col_h;
        if (!kFusion) {
          long long tw_w = 0, tw_x = 0;   // CN_TRACE_FINE: cycles this thread waits for weight slabs / X slabs in the first Dense
          // descriptors are built once per slab; a K step of 16 (32 bytes) is +2 on the encoded start address --
          // rebuilding them per MMA costs the issuing thread ~20 % of the tensor rate (tools/ubench_mma_issue.cu)
          uint32_t sw[6];
#pragma unroll
          for (int st = 0; st < 12; ++st) {
            const int kb = kL0StepKb[st], t = kL0StepT[st];
            if (t >= nt) continue;
            if (t == 0) {                                  // tile 0 is the first user of every weight slab
              const long long w0 = CN_TRACE_CLOCK_IF(true);
              sw[kb] = next_slot();
              tw_w += CN_TRACE_CLOCK_IF(true) - w0;
            }
            const uint64_t dw = umma_desc(slot_addr(sw[kb]), kKmajLbo, kSboBytes);
            const long long x0 = CN_TRACE_CLOCK_IF(true);
            const uint32_t sx = next_x_slot();
            const uint64_t dx = umma_desc(s_base + L::kXRing + sx * CN_SLAB_BYTES, kKmajLbo, kSboBytes);
            tw_x += CN_TRACE_CLOCK_IF(true) - x0;
            tc_fence_after();
#pragma unroll
            for (int j = 0; j < 4; ++j)
              umma_f16(tmem + t * 256 + col_h, dx + j * kKmajKstepEnc, dw + j * kKmajKstepEnc, idesc, (kb | j) ? 1u : 0u);
            umma_commit(smem_u32(&bars.xempty[sx]));
            if (kb == 5) umma_commit(smem_u32(&bars.h0_ready[t]));
            if (t == nt - 1) umma_commit(smem_u32(&bars.empty[sw[kb]]));   // last user of this weight slab
          }
          CN_TRACE_AT(pi, 1);
#ifdef CN_TRACE_FINE
          CN_TRACE_VAL(pi, 30, tw_w);
          CN_TRACE_VAL(pi, 31, tw_x);
#endif
        }
        for (int blk = 0; blk < kNumBlk; ++blk) {
          for (int j = 0; j < 2; ++j) {
            // ring order of the fusion stage: views(t0), Wa0, views(t1), Wb0, ...  The view slabs are read by
            // the epilogue warps, but this thread stays the ONLY in-order waiter on the ring's full barriers
            // (a waiter that skips ring rounds would alias mbarrier phase parities) and forwards "landed".
            if (kFusion && blk == 0 && j == 0 && view_slots) {
              for (uint32_t i = 0; i < view_slots; ++i) (void)next_slot();
              mbar_arrive(smem_u32(&bars.views_ready[0]));
            }
            const uint32_t s0 = next_slot();
            const uint32_t s1 = next_slot();
            for (int t = 0; t < nt; ++t) {
              if (kFusion && blk == 0 && j == 0 && t == 1 && view_slots) {
                for (uint32_t i = 0; i < view_slots; ++i) (void)next_slot();
                mbar_arrive(smem_u32(&bars.views_ready[1]));
              }
              const uint32_t act_parity = n_act[t] & 1u;
              ++n_act[t];
              if (blk == 0 && j == 0 && t == 0) CN_TRACE_AT(pi, 2);
              if (blk == kNumBlk - 1 && j == 1 && t == 1) CN_TRACE_AT(pi, 3);
              if (blk == 1 && t == 0) CN_TRACE_AT(pi, j ? 30 : 4);      // one steady-state layer of tile 0, end to end
              if (j == 0) {
                gemm_act(t, tmem + t * 256 + col_tmp, s0, s1, false, act_parity);
                umma_commit(smem_u32(&bars.tmp_ready[t]));
              } else {
                gemm_act(t, tmem + t * 256 + col_h, s0, s1, true, act_parity);
                umma_commit(smem_u32(&bars.h_ready[t]));
              }
              if (blk == 1 && t == 0) CN_TRACE_AT(pi, j ? 31 : 5);
            }
            umma_commit(smem_u32(&bars.empty[s0]));
            umma_commit(smem_u32(&bars.empty[s1]));
          }
        }
      }
    }
  } else if (warp < 18) {
    // =========================== epilogue: 8 warps per tile ===========================
    const uint32_t e = warp - 2;                    // 0..15
    const int t = e >> 3;                           // tile slot of this warp
    const uint32_t half = (e >> 2) & 1u;            // column half [64*half, 64*half + 64)
    const uint32_t quad = warp & 3u;                // TMEM lane quadrant this warp may access
    const uint32_t row = quad * 32u + lane;         // row of the tile == TMEM lane
    const bool wg_leader = (e & 7u) == 0 && lane == 0;
    const uint32_t bar_id = 1 + t;
    uint8_t *act = smem + kSmemAct + t * CN_ACT_TILE_BYTES;
    const uint32_t act_u32 = s_base + kSmemAct + t * CN_ACT_TILE_BYTES;
    const uint32_t t_base = tmem + ((quad * 32u) << 16) + t * 256;
    uint32_t n_h = 0, n_tmp = 0, n_views = 0;
    uint32_t ring_base = 0;                         // ring position at the start of the current pair
    bool store_pending = false;

    // make this thread's writes to ACT visible to the async proxy, optionally stream the tile to
    // global memory, then hand it to the MMA warp
    const uint32_t bar_s0 = kSlabSplit ? smem_u32(&bars.act_ready0[t]) : 0u;
    auto publish = [&](uint8_t *gdst, bool signal_mma, bool signal_slab0 = false) {
      fence_proxy_async_smem();
      if (gdst) {
        named_bar_sync(bar_id, kEpiThreads);
        if (wg_leader) {
          bulk_s2g(gdst, act_u32, CN_ACT_TILE_BYTES);
          bulk_commit();
        }
        store_pending = true;
      }
      if (signal_mma) {
        // 256 arrivals on one mbarrier serialise in shared memory: every lane fences, lane 0 arrives for the warp
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (kSlabSplit && signal_slab0) mbar_arrive(bar_s0);   // the writer covered the tile in one go (mean over views, chain transition)
          mbar_arrive(smem_u32(&bars.act_ready[t]));
        }
      }
    };
    // ACT may only be overwritten once a pending bulk store has finished reading it
    auto acquire = [&]() {
      if (store_pending) {
        if (wg_leader) bulk_wait_read0();
        named_bar_sync(bar_id, kEpiThreads);
        store_pending = false;
      }
    };

    int64_t pi = 0;
    const bool tr = threadIdx.x == 64;
    const bool tr1 = threadIdx.x == 64 + kEpiThreads;   // the same lane of tile slot 1 (coarse milestones only)
    for (int64_t p = blockIdx.x; p < n_pairs; p += gridDim.x, ++pi) {
      const int64_t tile = 2 * p + t;
      if (tile >= n_tiles) break;
      const uint32_t t_h = t_base + ((!kFusion && (pi & 1)) ? 128u : 0u);   // column roles alternate per pair (MMA issuer)
      const uint32_t t_tmp = t_base + ((!kFusion && (pi & 1)) ? 0u : 128u);
      uint8_t *my_acts = acts ? acts + tile * (int64_t)acts_stride * CN_ACT_TILE_BYTES : nullptr;
      const uint8_t *my_mask = mask_acts ? mask_acts + tile * (int64_t)acts_stride * CN_ACT_TILE_BYTES : nullptr;
      // fusion stage, pose-query models: taps[s][row][128] fp32, s = (mean, block 4, block 5, block 6 output)
      const int64_t tap_stride = n_tiles * (int64_t)(128 * 128);
      float *my_tap = ((kFusion || kChain) && taps && !tap_bf16) ? taps + (tile * 128 + row) * 128 : nullptr;
      // ... or as bf16 activation-tile images [4][n_tiles][32 KiB] (the operand tiles of cn_grasp_head_fwd)
      const int64_t tap16_stride = n_tiles * (int64_t)CN_ACT_TILE_BYTES;
      uint8_t *my_tap16 = ((kFusion || kChain) && taps && tap_bf16)
                              ? reinterpret_cast<uint8_t *>(taps) + tile * (int64_t)CN_ACT_TILE_BYTES : nullptr;

      if (kFusion) {
        // ---- mean over views -> fp32 residual stream in TMEM, relu(mean) -> first MMA operand
        const int64_t b = tile / tiles_per_view, tt = tile % tiles_per_view;
        const float inv_nv = 1.f / (float)nv;
        acquire();
        // ring position of this tile's view slabs: views(t0) | Wa0 (2 slots) | views(t1) | ...
        const uint32_t vbase = ring_base + (t ? view_slots + 2u : 0u);
        if (tr) CN_TRACE_AT(pi, 15);
        if (staged) {
          mbar_wait(smem_u32(&bars.views_ready[t]), n_views & 1u);
          ++n_views;
        }
        if (tr) CN_TRACE_AT(pi, 16);
        const uint8_t *vslab[kMaxStageViews];
#pragma unroll
        for (int v = 0; v < kMaxStageViews; ++v)
          vslab[v] = smem + kSmemRing + ((vbase + 2u * v + half) % kRingSlots) * CN_SLAB_BYTES;
#pragma unroll 1
        for (int c2 = 0; c2 < 2; ++c2) {
          uint32_t r[32];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t chunk = half * 8 + c2 * 4 + q;
            float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (staged) {
              // chunk's slab (= half) of view v sits in ring slot vbase + 2v + half.  All (<= 3) shared-memory
              // loads are issued before the first use: a per-view loop with a dependent accumulate after every
              // load made this prologue 4 900 cycles long (tools/trace_mlp_fusion.py).
              const uint32_t off = tile_chunk_off(row, chunk & 7u);
              uint4 x[kMaxStageViews];
#pragma unroll
              for (int v = 0; v < kMaxStageViews; ++v)
                x[v] = v < nv ? *reinterpret_cast<const uint4 *>(vslab[v] + off) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
              for (int v = 0; v < kMaxStageViews; ++v) {
                acc[0] += bf16_lo(x[v].x); acc[1] += bf16_hi(x[v].x); acc[2] += bf16_lo(x[v].y); acc[3] += bf16_hi(x[v].y);
                acc[4] += bf16_lo(x[v].z); acc[5] += bf16_hi(x[v].z); acc[6] += bf16_lo(x[v].w); acc[7] += bf16_hi(x[v].w);
              }
            } else {
              for (int v = 0; v < nv; ++v) {
                const uint4 x = ldg_nc_u4(in_tiles + ((b * nv + v) * tiles_per_view + tt) * (int64_t)CN_ACT_TILE_BYTES +
                                          tile_chunk_off(row, chunk));
                acc[0] += bf16_lo(x.x); acc[1] += bf16_hi(x.x); acc[2] += bf16_lo(x.y); acc[3] += bf16_hi(x.y);
                acc[4] += bf16_lo(x.z); acc[5] += bf16_hi(x.z); acc[6] += bf16_lo(x.w); acc[7] += bf16_hi(x.w);
              }
            }
            float m[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              m[i] = acc[i] * inv_nv;
              r[q * 8 + i] = __float_as_uint(m[i]);
            }
            if (my_tap) {
              *reinterpret_cast<float4 *>(my_tap + chunk * 8) = make_float4(m[0], m[1], m[2], m[3]);
              *reinterpret_cast<float4 *>(my_tap + chunk * 8 + 4) = make_float4(m[4], m[5], m[6], m[7]);
            }
            if (my_tap16)
              *reinterpret_cast<uint4 *>(my_tap16 + tile_chunk_off(row, chunk)) =
                  make_uint4(pack_bf16x2(m[0], m[1]), pack_bf16x2(m[2], m[3]), pack_bf16x2(m[4], m[5]),
                             pack_bf16x2(m[6], m[7]));
            if (my_mask) {
              const uint4 mk = ldg_nc_u4(my_mask + tile_chunk_off(row, chunk));
              *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, chunk)) =
                  make_uint4(mul_bf16x2(pack_bf16x2(m[0], m[1]), gt0_bf16x2(mk.x)), mul_bf16x2(pack_bf16x2(m[2], m[3]), gt0_bf16x2(mk.y)),
                             mul_bf16x2(pack_bf16x2(m[4], m[5]), gt0_bf16x2(mk.z)), mul_bf16x2(pack_bf16x2(m[6], m[7]), gt0_bf16x2(mk.w)));
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i) m[i] = fmaxf(m[i], 0.f);
              *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, chunk)) =
                  make_uint4(pack_bf16x2(m[0], m[1]), pack_bf16x2(m[2], m[3]), pack_bf16x2(m[4], m[5]),
                             pack_bf16x2(m[6], m[7]));
            }
          }
          tmem_st32(t_h + half * 64 + c2 * 32, r);
        }
        if (tr) CN_TRACE_AT(pi, 19);
        if (staged) {
          // every thread has read its rows: hand the view slots back to the producer
          named_bar_sync(bar_id, kEpiThreads);
          if (wg_leader)
            for (uint32_t i = 0; i < view_slots; ++i) mbar_arrive(smem_u32(&bars.empty[(vbase + i) % kRingSlots]));
        }
        tmem_wait_st();
        if (tr) CN_TRACE_AT(pi, 20);
        publish(my_acts, true, true);
        if (tr) CN_TRACE_AT(pi, 17);
      } else {
        mbar_wait(smem_u32(&bars.h0_ready[t]), (uint32_t)pi & 1u);     // one completion per pair
        tc_fence_after();
        if (tr) CN_TRACE_AT(pi, 8);
      }

      // one ResNet block of a stage: tb = the stage's cumulative-bias tables, sacts = its saved-activation tiles of this
      // tile (n_acts regular ones, the hidden activations behind them), smask = the tangent pass's mask tiles
      long long tr_epi = 0, tr_wait_tmp = 0, tr_wait_h = 0, tr_pub = 0;   // CN_TRACE: cycles of thread `tr` in epilogues / waiting for the MMAs
      auto run_block = [&](int blk, bool skip_ea, const float *tb, uint8_t *sacts, int n_acts, const uint8_t *smask,
                           bool with_taps) {
        long long c0 = CN_TRACE_CLOCK_IF(tr);
        if (!skip_ea) {
          // E_a: relu(h + cumulative bias) -> operand of the block's first Dense
          acquire();
          epi_tmem_to_act<true, kPhased>(t_h, tb + (2 * blk) * 128, act, row, half, bar_s0,
                                (with_taps && my_tap) ? my_tap + blk * tap_stride : nullptr,
                                (with_taps && my_tap16) ? my_tap16 + blk * tap16_stride : nullptr,
                                smask ? smask + blk * CN_ACT_TILE_BYTES : nullptr);
          const long long p0 = CN_TRACE_CLOCK_IF(tr);
          publish(sacts ? sacts + blk * CN_ACT_TILE_BYTES : nullptr, true);
          tr_pub += CN_TRACE_CLOCK_IF(tr) - p0;
          if (tr && blk == 0) CN_TRACE_AT(pi, 9);
        }
        // E_b: relu(t + ba) -> operand of the block's second Dense
        long long c1 = CN_TRACE_CLOCK_IF(tr);
        mbar_wait(smem_u32(&bars.tmp_ready[t]), n_tmp & 1u);
        ++n_tmp;
        tc_fence_after();
        long long c2 = CN_TRACE_CLOCK_IF(tr);
        if (tr && blk == 0) CN_TRACE_AT(pi, 10);
        if (tr && blk == 1 && tb == tab) CN_TRACE_AT(pi, 6);
        acquire();
        epi_tmem_to_act<true, kPhased>(t_tmp, tb + (2 * blk + 1) * 128, act, row, half, bar_s0, nullptr, nullptr,
                              smask ? smask + (n_acts + blk) * CN_ACT_TILE_BYTES : nullptr);
        if (tr && blk == 1 && tb == tab) CN_TRACE_AT(pi, 28);
        const long long p1 = CN_TRACE_CLOCK_IF(tr);
        publish((save_hidden && sacts) ? sacts + (n_acts + blk) * CN_ACT_TILE_BYTES : nullptr, true);
        tr_pub += CN_TRACE_CLOCK_IF(tr) - p1;
        if (tr && blk == 1 && tb == tab) CN_TRACE_AT(pi, 29);
        if (tr && blk == 0) CN_TRACE_AT(pi, 11);
        long long c3 = CN_TRACE_CLOCK_IF(tr);
        mbar_wait(smem_u32(&bars.h_ready[t]), n_h & 1u);
        ++n_h;
        tc_fence_after();
        if (tr && blk == 0) CN_TRACE_AT(pi, 12);
        if (tr && blk == 1 && tb == tab) CN_TRACE_AT(pi, 7);
        long long c4 = CN_TRACE_CLOCK_IF(tr);
        tr_epi += (c1 - c0) + (c3 - c2);
        tr_wait_tmp += c2 - c1;
        tr_wait_h += c4 - c3;
      };
      auto readout = [&](const float *tb, uint8_t *sacts) {
        // ---- readout: relu(e) . Wr + br, sigmoid / softplus (layers.py:392-397).  The two column-half
        // warps of a row split the OUTPUTS (half 0: r,g; half 1: b,sigma), each over all 128 inputs; the
        // relu(e) tile saved for the backward is still written half by half.
        const float *wr = tab + kRoOff;
        // two accumulator chains per output: one chain of 128 dependent FMAs per output made this epilogue latency bound
        float ya = wr[512 + 2 * half], yb = wr[512 + 2 * half + 1], ya1 = 0.f, yb1 = 0.f;
#pragma unroll 1
        for (int c4 = 0; c4 < 4; ++c4) {
          uint32_t r[32];
          tmem_ld32(t_h + c4 * 32, r);
          tmem_wait_ld();
          const bool mine = (uint32_t)(c4 >> 1) == half;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int col = c4 * 32 + q * 8 + i;
              v[i] = fmaxf(__uint_as_float(r[q * 8 + i]) + tb[6 * 128 + col], 0.f);
              const float2 w = *reinterpret_cast<const float2 *>(wr + col * 4 + 2 * half);
              if (i & 1) {
                ya1 += v[i] * w.x;
                yb1 += v[i] * w.y;
              } else {
                ya += v[i] * w.x;
                yb += v[i] * w.y;
              }
            }
            if ((my_tap || my_tap16) && mine) {
              // e itself (pre-ReLU) is the tap; v[] holds relu(e), so re-add from the accumulator
              float ev[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) ev[i] = __uint_as_float(r[q * 8 + i]) + tb[6 * 128 + c4 * 32 + q * 8 + i];
              if (my_tap) {
                float *dst = my_tap + 3 * tap_stride + c4 * 32 + q * 8;
                *reinterpret_cast<float4 *>(dst) = make_float4(ev[0], ev[1], ev[2], ev[3]);
                *reinterpret_cast<float4 *>(dst + 4) = make_float4(ev[4], ev[5], ev[6], ev[7]);
              } else {
                *reinterpret_cast<uint4 *>(my_tap16 + 3 * tap16_stride + tile_chunk_off(row, c4 * 4 + q)) =
                    make_uint4(pack_bf16x2(ev[0], ev[1]), pack_bf16x2(ev[2], ev[3]), pack_bf16x2(ev[4], ev[5]),
                               pack_bf16x2(ev[6], ev[7]));
              }
            }
            if (sacts && mine)
              *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, c4 * 4 + q)) =
                  make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                             pack_bf16x2(v[6], v[7]));
          }
        }
        ya += ya1;
        yb += yb1;
        float2 o;
        o.x = 1.f / (1.f + expf(-ya));
        o.y = half ? (yb > 20.f ? yb : log1pf(expf(yb))) : 1.f / (1.f + expf(-yb));
        reinterpret_cast<float2 *>(rgbs)[(tile * 128 + row) * 2 + half] = o;
        publish(sacts ? sacts + 3 * CN_ACT_TILE_BYTES : nullptr, false);
      };

      for (int blk = 0; blk < 3; ++blk) run_block(blk, kFusion && blk == 0, tab, my_acts, kActsPerTile, my_mask, kFusion);
      if (tr) CN_TRACE_AT(pi, 13);
      if (tr1) CN_TRACE_AT(pi, 18);

      acquire();
      if (kFusion) {
        readout(tab, my_acts);
      } else if (!kChain) {
        // ---- h3 = h + cumulative bias -> bf16 tile for the fusion stage
        epi_tmem_to_act<false>(t_h, tab + 6 * 128, act, row, half, 0u);
        publish(out_tiles + tile * (int64_t)CN_ACT_TILE_BYTES, false);
      } else {
        // ---- chained fusion stage (nv = 1).  m = bf16(h + cumulative bias) is what the two-kernel path stores as the
        // per-view output and reads back as the mean over its one view: the fp32 stream restarts from it.
        const float *tabu = tab + 7 * 128;
        const int acts_stride_u = 4 + ((save_hidden || mask_acts_u) ? 3 : 0);
        uint8_t *my_acts_u = acts_u ? acts_u + tile * (int64_t)acts_stride_u * CN_ACT_TILE_BYTES : nullptr;
        const uint8_t *my_mask_u = mask_acts_u ? mask_acts_u + tile * (int64_t)acts_stride_u * CN_ACT_TILE_BYTES : nullptr;
#pragma unroll 1
        for (int c2 = 0; c2 < 2; ++c2) {
          uint32_t r[32];
          tmem_ld32(t_h + half * 64 + c2 * 32, r);
          tmem_wait_ld();
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t chunk = half * 8 + c2 * 4 + q;
            const float *bq = tab + 6 * 128 + chunk * 8;
            uint32_t w[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              w[i] = cvt_bf16x2(__uint_as_float(r[q * 8 + 2 * i]) + bq[2 * i], __uint_as_float(r[q * 8 + 2 * i + 1]) + bq[2 * i + 1]);
              r[q * 8 + 2 * i] = __float_as_uint(bf16_lo(w[i]));
              r[q * 8 + 2 * i + 1] = __float_as_uint(bf16_hi(w[i]));
            }
            if (my_tap) {
              *reinterpret_cast<float4 *>(my_tap + chunk * 8) = make_float4(__uint_as_float(r[q * 8]), __uint_as_float(r[q * 8 + 1]),
                                                                            __uint_as_float(r[q * 8 + 2]), __uint_as_float(r[q * 8 + 3]));
              *reinterpret_cast<float4 *>(my_tap + chunk * 8 + 4) = make_float4(__uint_as_float(r[q * 8 + 4]), __uint_as_float(r[q * 8 + 5]),
                                                                                __uint_as_float(r[q * 8 + 6]), __uint_as_float(r[q * 8 + 7]));
            }
            if (my_tap16) *reinterpret_cast<uint4 *>(my_tap16 + tile_chunk_off(row, chunk)) = make_uint4(w[0], w[1], w[2], w[3]);
            uint4 o;
            if (my_mask_u) {
              const uint4 mk = ldg_nc_u4(my_mask_u + tile_chunk_off(row, chunk));
              o = make_uint4(mul_bf16x2(w[0], gt0_bf16x2(mk.x)), mul_bf16x2(w[1], gt0_bf16x2(mk.y)),
                             mul_bf16x2(w[2], gt0_bf16x2(mk.z)), mul_bf16x2(w[3], gt0_bf16x2(mk.w)));
            } else {
              float m[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) m[i] = fmaxf(__uint_as_float(r[q * 8 + i]), 0.f);
              o = make_uint4(pack_bf16x2(m[0], m[1]), pack_bf16x2(m[2], m[3]), pack_bf16x2(m[4], m[5]), pack_bf16x2(m[6], m[7]));
            }
            *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, chunk)) = o;
          }
          tmem_st32(t_h + half * 64 + c2 * 32, r);
        }
        tmem_wait_st();
        publish(my_acts_u, true, true);
        for (int blk = 0; blk < 3; ++blk) run_block(blk, blk == 0, tabu, my_acts_u, 4, my_mask_u, true);
        acquire();
        readout(tabu, my_acts_u);
      }
      if (tr) CN_TRACE_AT(pi, 14);
      if (tr1) CN_TRACE_AT(pi, 23);
      if (tr) {
        CN_TRACE_VAL(pi, 24, tr_epi);
        CN_TRACE_VAL(pi, 25, tr_wait_tmp);
        CN_TRACE_VAL(pi, 26, tr_wait_h);
        CN_TRACE_VAL(pi, 27, tr_pub);
      }
      ring_base += 12u + view_slots * ((2 * p + 1 < n_tiles) ? 2u : 1u);
    }
    if (wg_leader) bulk_wait0();
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

CN_TRACE_READER(cn_debug_read_trace_fwd)

template <bool kFusion, bool kExt, bool kChain = false>
static cudaError_t launch_fwd_t(unsigned grid, const void *packed, const void *in_tiles, int64_t n_tiles, int nv,
                                int64_t tiles_per_view, void *out_tiles, float *rgbs, void *acts, float *taps,
                                cudaStream_t st, int tap_bf16, int save_hidden, const void *mask_acts,
                                void *acts_u = nullptr, const void *mask_acts_u = nullptr) {
  constexpr uint32_t bytes = FwdSmem<kFusion, kChain>::kBytes;
  cudaError_t e = cudaFuncSetAttribute(mlp_fwd_kernel<kFusion, kExt, kChain>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)bytes);
  if (e == cudaSuccess)
    mlp_fwd_kernel<kFusion, kExt, kChain><<<grid, kFwdThreads, bytes, st>>>(
        (const uint8_t *)packed, (const uint8_t *)in_tiles, n_tiles, nv, tiles_per_view, (uint8_t *)out_tiles, rgbs,
        (uint8_t *)acts, taps, tap_bf16, save_hidden, (const uint8_t *)mask_acts, (uint8_t *)acts_u,
        (const uint8_t *)mask_acts_u);
  return e;
}

static int launch_fwd(bool fusion, const void *packed, const void *in_tiles, int64_t n_tiles, int nv,
                      int64_t tiles_per_view, void *out_tiles, float *rgbs, void *acts, float *taps,
                      cudaStream_t st, int tap_bf16 = 0, int save_hidden = 0, const void *mask_acts = nullptr) {
  const int64_t n_pairs = (n_tiles + 1) / 2;
  const unsigned grid = (unsigned)(n_pairs < num_sms() ? n_pairs : num_sms());
  const bool ext = tap_bf16 || save_hidden || mask_acts;
  cudaError_t e;
  if (fusion)
    e = ext ? launch_fwd_t<true, true>(grid, packed, in_tiles, n_tiles, nv, tiles_per_view, out_tiles, rgbs, acts, taps, st,
                                       tap_bf16, save_hidden, mask_acts)
            : launch_fwd_t<true, false>(grid, packed, in_tiles, n_tiles, nv, tiles_per_view, out_tiles, rgbs, acts, taps, st,
                                        0, 0, nullptr);
  else
    e = ext ? launch_fwd_t<false, true>(grid, packed, in_tiles, n_tiles, nv, tiles_per_view, out_tiles, rgbs, acts, nullptr,
                                        st, 0, save_hidden, mask_acts)
            : launch_fwd_t<false, false>(grid, packed, in_tiles, n_tiles, nv, tiles_per_view, out_tiles, rgbs, acts, nullptr,
                                         st, 0, 0, nullptr);
  if (e != cudaSuccess) {
    set_error("mlp_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return check_launch(fusion ? "cn_mlp_fusion_fwd" : "cn_mlp_feature_fwd");
}

}  // namespace cn

extern "C" int cn_mlp_feature_fwd(const void *packed, const void *x_tiles, int64_t n_tiles, void *h3_tiles, void *acts,
                                  void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && x_tiles && h3_tiles, "cn_mlp_feature_fwd: null pointer");
  CN_REQUIRE(n_tiles >= 0, "cn_mlp_feature_fwd: n_tiles %lld", (long long)n_tiles);
  CN_REQUIRE((((uintptr_t)packed | (uintptr_t)x_tiles | (uintptr_t)h3_tiles | (uintptr_t)acts) & 127) == 0,
             "cn_mlp_feature_fwd: buffers must be 128-byte aligned");
  if (n_tiles == 0) return 0;
  return launch_fwd(false, packed, x_tiles, n_tiles, 1, n_tiles, h3_tiles, nullptr, acts, nullptr,
                    (cudaStream_t)stream);
}

extern "C" int cn_mlp_fusion_fwd(const void *packed, const void *h3_tiles, int32_t b, int32_t nv,
                                 int64_t tiles_per_view, float *rgbs, void *acts, void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && h3_tiles && rgbs, "cn_mlp_fusion_fwd: null pointer");
  CN_REQUIRE(b > 0 && nv > 0 && tiles_per_view >= 0, "cn_mlp_fusion_fwd: b=%d nv=%d tiles_per_view=%lld", b, nv,
             (long long)tiles_per_view);
  CN_REQUIRE((((uintptr_t)packed | (uintptr_t)h3_tiles | (uintptr_t)rgbs | (uintptr_t)acts) & 127) == 0,
             "cn_mlp_fusion_fwd: buffers must be 128-byte aligned");
  if (tiles_per_view == 0) return 0;
  return launch_fwd(true, packed, h3_tiles, (int64_t)b * tiles_per_view, nv, tiles_per_view, nullptr, rgbs, acts,
                    nullptr, (cudaStream_t)stream);
}

extern "C" int cn_mlp_fusion_fwd_taps(const void *packed, const void *h3_tiles, int32_t b, int32_t nv,
                                      int64_t tiles_per_view, float *rgbs, void *acts, float *taps, void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && h3_tiles && rgbs && taps, "cn_mlp_fusion_fwd_taps: null pointer");
  CN_REQUIRE(b > 0 && nv > 0 && tiles_per_view >= 0, "cn_mlp_fusion_fwd_taps: b=%d nv=%d tiles_per_view=%lld", b, nv,
             (long long)tiles_per_view);
  CN_REQUIRE((((uintptr_t)packed | (uintptr_t)h3_tiles | (uintptr_t)rgbs | (uintptr_t)acts | (uintptr_t)taps) & 127) == 0,
             "cn_mlp_fusion_fwd_taps: buffers must be 128-byte aligned");
  if (tiles_per_view == 0) return 0;
  return launch_fwd(true, packed, h3_tiles, (int64_t)b * tiles_per_view, nv, tiles_per_view, nullptr, rgbs, acts,
                    taps, (cudaStream_t)stream);
}

extern "C" int cn_mlp_fusion_fwd_tap_tiles(const void *packed, const void *h3_tiles, int32_t b, int32_t nv,
                                           int64_t tiles_per_view, float *rgbs, void *acts, void *tap_tiles, void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && h3_tiles && rgbs && tap_tiles, "cn_mlp_fusion_fwd_tap_tiles: null pointer");
  CN_REQUIRE(b > 0 && nv > 0 && tiles_per_view >= 0, "cn_mlp_fusion_fwd_tap_tiles: b=%d nv=%d tiles_per_view=%lld", b, nv,
             (long long)tiles_per_view);
  CN_REQUIRE((((uintptr_t)packed | (uintptr_t)h3_tiles | (uintptr_t)rgbs | (uintptr_t)acts | (uintptr_t)tap_tiles) & 127) == 0,
             "cn_mlp_fusion_fwd_tap_tiles: buffers must be 128-byte aligned");
  if (tiles_per_view == 0) return 0;
  return launch_fwd(true, packed, h3_tiles, (int64_t)b * tiles_per_view, nv, tiles_per_view, nullptr, rgbs, acts,
                    (float *)tap_tiles, (cudaStream_t)stream, 1);
}

// ---- pose-query training (delta_ngf/model.py:264-296: d/dW of a function of d prediction / d pose)
extern "C" int cn_mlp_feature_fwd_ex(const void *packed, const void *x_tiles, int64_t n_tiles, void *h3_tiles, void *acts,
                                     int32_t save_hidden, const void *mask_acts, void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && x_tiles && h3_tiles, "cn_mlp_feature_fwd_ex: null pointer");
  CN_REQUIRE(n_tiles >= 0, "cn_mlp_feature_fwd_ex: n_tiles %lld", (long long)n_tiles);
  CN_REQUIRE(!(mask_acts && acts), "cn_mlp_feature_fwd_ex: the tangent pass (mask_acts) saves nothing");
  CN_REQUIRE((((uintptr_t)packed | (uintptr_t)x_tiles | (uintptr_t)h3_tiles | (uintptr_t)acts | (uintptr_t)mask_acts) & 127) == 0,
             "cn_mlp_feature_fwd_ex: buffers must be 128-byte aligned");
  if (n_tiles == 0) return 0;
  return launch_fwd(false, packed, x_tiles, n_tiles, 1, n_tiles, h3_tiles, nullptr, acts, nullptr, (cudaStream_t)stream, 0,
                    save_hidden, mask_acts);
}

extern "C" int cn_mlp_fusion_fwd_taps_ex(const void *packed, const void *h3_tiles, int32_t b, int32_t nv,
                                         int64_t tiles_per_view, float *rgbs, void *acts, void *taps, int32_t tap_bf16,
                                         int32_t save_hidden, const void *mask_acts, void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && h3_tiles && rgbs && taps, "cn_mlp_fusion_fwd_taps_ex: null pointer");
  CN_REQUIRE(b > 0 && nv > 0 && tiles_per_view >= 0, "cn_mlp_fusion_fwd_taps_ex: b=%d nv=%d tiles_per_view=%lld", b, nv,
             (long long)tiles_per_view);
  CN_REQUIRE(!(mask_acts && acts), "cn_mlp_fusion_fwd_taps_ex: the tangent pass (mask_acts) saves nothing");
  CN_REQUIRE((((uintptr_t)packed | (uintptr_t)h3_tiles | (uintptr_t)rgbs | (uintptr_t)acts | (uintptr_t)taps |
               (uintptr_t)mask_acts) & 127) == 0, "cn_mlp_fusion_fwd_taps_ex: buffers must be 128-byte aligned");
  if (tiles_per_view == 0) return 0;
  return launch_fwd(true, packed, h3_tiles, (int64_t)b * tiles_per_view, nv, tiles_per_view, nullptr, rgbs, acts,
                    (float *)taps, (cudaStream_t)stream, tap_bf16, save_hidden, mask_acts);
}

// ---- nv = 1: per-view stage + fusion stage + readout in one launch (the mean over one view is the identity)
extern "C" int cn_mlp_chain_fwd(const void *packed, const void *x_tiles, int64_t n_tiles, float *rgbs, void *acts_f,
                                void *acts_u, void *taps, int32_t tap_mode, int32_t save_hidden, const void *mask_acts_f,
                                const void *mask_acts_u, void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && x_tiles && rgbs, "cn_mlp_chain_fwd: null pointer");
  CN_REQUIRE(n_tiles >= 0, "cn_mlp_chain_fwd: n_tiles %lld", (long long)n_tiles);
  CN_REQUIRE(tap_mode >= 0 && tap_mode <= 2 && (tap_mode == 0) == (taps == nullptr), "cn_mlp_chain_fwd: tap_mode %d (0 none, "
             "1 fp32 rows, 2 bf16 tiles) does not match taps", tap_mode);
  CN_REQUIRE((acts_f == nullptr) == (acts_u == nullptr), "cn_mlp_chain_fwd: acts_f and acts_u go together");
  CN_REQUIRE((mask_acts_f == nullptr) == (mask_acts_u == nullptr) && !(mask_acts_f && acts_f),
             "cn_mlp_chain_fwd: the tangent pass takes both mask buffers and saves nothing");
  CN_REQUIRE((((uintptr_t)packed | (uintptr_t)x_tiles | (uintptr_t)rgbs | (uintptr_t)acts_f | (uintptr_t)acts_u | (uintptr_t)taps |
               (uintptr_t)mask_acts_f | (uintptr_t)mask_acts_u) & 127) == 0, "cn_mlp_chain_fwd: buffers must be 128-byte aligned");
  if (n_tiles == 0) return 0;
  const int64_t n_pairs = (n_tiles + 1) / 2;
  const unsigned grid = (unsigned)(n_pairs < num_sms() ? n_pairs : num_sms());
  const bool ext = tap_mode || save_hidden || mask_acts_f;
  cudaError_t e = ext ? launch_fwd_t<false, true, true>(grid, packed, x_tiles, n_tiles, 1, n_tiles, nullptr, rgbs, acts_f,
                                                        (float *)taps, (cudaStream_t)stream, tap_mode == 2, save_hidden,
                                                        mask_acts_f, acts_u, mask_acts_u)
                      : launch_fwd_t<false, false, true>(grid, packed, x_tiles, n_tiles, 1, n_tiles, nullptr, rgbs, acts_f,
                                                         nullptr, (cudaStream_t)stream, 0, 0, nullptr, acts_u, nullptr);
  if (e != cudaSuccess) {
    set_error("cn_mlp_chain_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return check_launch("cn_mlp_chain_fwd");
}
