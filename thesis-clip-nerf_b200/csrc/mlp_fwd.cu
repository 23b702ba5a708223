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
//   warps 2-5   epilogue tile 0: TMEM -> registers -> bias / ReLU -> bf16 -> swizzled smem operand
//   warps 6-9   epilogue tile 1
// TMEM (512 columns): per tile a 128-column fp32 residual stream `h` and a 128-column `tmp`.
// The second GEMM of every block accumulates straight into `h`, so the residual add is free and
// the stream stays fp32; Dense biases of the residual branch are applied as cumulative bias
// vectors when `h` is read (tables built by cn_pack_weights).
#include "common.cuh"
#include "mlp_layout.cuh"

namespace cn {

constexpr int kRingSlots = 8;
constexpr int kFwdThreads = 320;
constexpr uint32_t kSmemAct = 0;                                   // 2 x 32 KiB
constexpr uint32_t kSmemRing = 2 * CN_ACT_TILE_BYTES;              // 8 x 16 KiB
constexpr uint32_t kSmemTab = kSmemRing + kRingSlots * CN_SLAB_BYTES;
constexpr uint32_t kSmemTabBytes = 7 * 128 * 4 + 512 * 4 + 16;
constexpr uint32_t kFwdSmemBytes = kSmemTab + kSmemTabBytes + 1024;  // + alignment slack

struct FwdBars {
  uint64_t full[kRingSlots];
  uint64_t empty[kRingSlots];
  uint64_t h_ready[2];
  uint64_t tmp_ready[2];
  uint64_t act_ready[2];
  uint64_t h_free[2];
};

// TMEM accumulator (this thread's lane, 128 columns) + bias -> (ReLU) -> bf16 row of the
// swizzled activation tile in shared memory.
template <bool kRelu>
__device__ __forceinline__ void epi_tmem_to_act(uint32_t taddr, const float *__restrict__ bias, uint8_t *act,
                                                uint32_t row) {
#pragma unroll 1
  for (int c4 = 0; c4 < 4; ++c4) {
    uint32_t r[32];
    tmem_ld32(taddr + c4 * 32, r);
    tmem_wait_ld();
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 b0 = *reinterpret_cast<const float4 *>(bias + c4 * 32 + q * 8);
      const float4 b1 = *reinterpret_cast<const float4 *>(bias + c4 * 32 + q * 8 + 4);
      float v[8];
      v[0] = __uint_as_float(r[q * 8 + 0]) + b0.x; v[1] = __uint_as_float(r[q * 8 + 1]) + b0.y;
      v[2] = __uint_as_float(r[q * 8 + 2]) + b0.z; v[3] = __uint_as_float(r[q * 8 + 3]) + b0.w;
      v[4] = __uint_as_float(r[q * 8 + 4]) + b1.x; v[5] = __uint_as_float(r[q * 8 + 5]) + b1.y;
      v[6] = __uint_as_float(r[q * 8 + 6]) + b1.z; v[7] = __uint_as_float(r[q * 8 + 7]) + b1.w;
      if (kRelu) {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
      }
      const uint4 o = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                 pack_bf16x2(v[6], v[7]));
      *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, c4 * 4 + q)) = o;
    }
  }
}

template <bool kFusion>
__global__ void __launch_bounds__(kFwdThreads, 1)
mlp_fwd_kernel(const uint8_t *__restrict__ packed, const uint8_t *__restrict__ in_tiles, int64_t n_tiles,
               int nv, int64_t tiles_per_view, uint8_t *__restrict__ out_tiles, float *__restrict__ rgbs,
               uint8_t *__restrict__ acts) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) FwdBars bars;
  __shared__ uint32_t tmem_slot;

  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t s_base = smem_u32(smem);
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  float *tab = reinterpret_cast<float *>(smem + kSmemTab);
  constexpr int kActsPerTile = kFusion ? 7 : 6;
  constexpr int kBlk0 = kFusion ? 3 : 0;

  // ---- one-time setup
  {
    const float *gt = reinterpret_cast<const float *>(packed + kPackTables);
    for (int i = threadIdx.x; i < 7 * 128; i += kFwdThreads) tab[i] = gt[(kFusion ? kTabFusion : kTabFeature) + i];
    if (kFusion)
      for (int i = threadIdx.x; i < 516; i += kFwdThreads) tab[7 * 128 + i] = gt[kTabReadoutW + i];
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < kRingSlots; ++i) {
      mbar_init(smem_u32(&bars.full[i]), 1);
      mbar_init(smem_u32(&bars.empty[i]), 1);
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(smem_u32(&bars.h_ready[t]), 1);
      mbar_init(smem_u32(&bars.tmp_ready[t]), 1);
      mbar_init(smem_u32(&bars.act_ready[t]), 128);
      mbar_init(smem_u32(&bars.h_free[t]), 128);
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
      for (int64_t p = blockIdx.x; p < n_pairs; p += gridDim.x) {
        const bool valid1 = 2 * p + 1 < n_tiles;
        if (!kFusion) {
          for (int kb = 0; kb < 6; ++kb) {
            load(packed + kPackFwdW0 + kb * CN_SLAB_BYTES);
            load(in_tiles + (2 * p) * (int64_t)CN_X_TILE_BYTES + kb * CN_SLAB_BYTES);
            if (valid1) load(in_tiles + (2 * p + 1) * (int64_t)CN_X_TILE_BYTES + kb * CN_SLAB_BYTES);
          }
        }
        for (int blk = 0; blk < 3; ++blk)
          for (int j = 0; j < 2; ++j)
            for (int s = 0; s < 2; ++s) load(packed + pack_fwd_block(kBlk0 + blk, j) + s * CN_SLAB_BYTES);
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(128, 128, 0, 0);
      uint32_t cc = 0;
      uint32_t n_act[2] = {0, 0}, n_free[2] = {0, 0};
      auto next_slot = [&]() -> uint32_t {
        const uint32_t slot = cc % kRingSlots;
        mbar_wait(smem_u32(&bars.full[slot]), (cc / kRingSlots) & 1u);
        ++cc;
        return slot;
      };
      auto slot_addr = [&](uint32_t slot) { return s_base + kSmemRing + slot * CN_SLAB_BYTES; };
      // D(tmem) (+)= ACT[t] (2 slabs) . W (2 ring slots)
      auto gemm_act = [&](int t, uint32_t d_tmem, uint32_t s0, uint32_t s1, bool accumulate) {
        const uint32_t a_base = s_base + kSmemAct + t * CN_ACT_TILE_BYTES;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const uint32_t b_base = slot_addr(s ? s1 : s0);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            umma_f16(d_tmem, umma_desc(a_base + s * CN_SLAB_BYTES + j * kKmajKstep, kKmajLbo, kSboBytes),
                     umma_desc(b_base + j * kKmajKstep, kKmajLbo, kSboBytes), idesc,
                     (accumulate || s || j) ? 1u : 0u);
        }
      };
      for (int64_t p = blockIdx.x; p < n_pairs; p += gridDim.x) {
        const int nt = (2 * p + 1 < n_tiles) ? 2 : 1;
        if (!kFusion) {
          // the previous pair's final epilogue must have drained h before it is overwritten
          for (int t = 0; t < nt; ++t) {
            mbar_wait(smem_u32(&bars.h_free[t]), (n_free[t] & 1u) ^ 1u);
            ++n_free[t];
          }
          tc_fence_after();
          for (int kb = 0; kb < 6; ++kb) {
            const uint32_t sw = next_slot();
            for (int t = 0; t < nt; ++t) {
              const uint32_t sx = next_slot();
              tc_fence_after();
#pragma unroll
              for (int j = 0; j < 4; ++j)
                umma_f16(tmem + t * 256, umma_desc(slot_addr(sx) + j * kKmajKstep, kKmajLbo, kSboBytes),
                         umma_desc(slot_addr(sw) + j * kKmajKstep, kKmajLbo, kSboBytes), idesc, (kb | j) ? 1u : 0u);
              umma_commit(smem_u32(&bars.empty[sx]));
              if (kb == 5) umma_commit(smem_u32(&bars.h_ready[t]));
            }
            umma_commit(smem_u32(&bars.empty[sw]));
          }
        }
        for (int blk = 0; blk < 3; ++blk) {
          for (int j = 0; j < 2; ++j) {
            const uint32_t s0 = next_slot();
            const uint32_t s1 = next_slot();
            for (int t = 0; t < nt; ++t) {
              mbar_wait(smem_u32(&bars.act_ready[t]), n_act[t] & 1u);
              ++n_act[t];
              tc_fence_after();
              if (j == 0) {
                gemm_act(t, tmem + t * 256 + 128, s0, s1, false);
                umma_commit(smem_u32(&bars.tmp_ready[t]));
              } else {
                gemm_act(t, tmem + t * 256, s0, s1, true);
                umma_commit(smem_u32(&bars.h_ready[t]));
              }
            }
            umma_commit(smem_u32(&bars.empty[s0]));
            umma_commit(smem_u32(&bars.empty[s1]));
          }
        }
      }
    }
  } else {
    // =========================== epilogue warpgroups ===========================
    const int t = (warp - 2) >> 2;                  // tile slot of this warpgroup
    const uint32_t quad = warp & 3u;                // TMEM lane quadrant this warp may access
    const uint32_t row = quad * 32u + lane;         // row of the tile == TMEM lane
    const bool wg_leader = ((warp - 2) & 3u) == 0 && lane == 0;
    const uint32_t bar_id = 1 + t;
    uint8_t *act = smem + kSmemAct + t * CN_ACT_TILE_BYTES;
    const uint32_t act_u32 = s_base + kSmemAct + t * CN_ACT_TILE_BYTES;
    const uint32_t t_h = tmem + ((quad * 32u) << 16) + t * 256;
    const uint32_t t_tmp = t_h + 128;
    uint32_t n_h = 0, n_tmp = 0;
    bool store_pending = false;

    // make this thread's writes to ACT visible to the async proxy, optionally stream the tile to
    // global memory, then hand it to the MMA warp
    auto publish = [&](uint8_t *gdst, bool signal_mma) {
      fence_proxy_async_smem();
      if (gdst) {
        named_bar_sync(bar_id, 128);
        if (wg_leader) {
          bulk_s2g(gdst, act_u32, CN_ACT_TILE_BYTES);
          bulk_commit();
        }
        store_pending = true;
      }
      if (signal_mma) {
        tc_fence_before();
        mbar_arrive(smem_u32(&bars.act_ready[t]));
      }
    };
    // ACT may only be overwritten once a pending bulk store has finished reading it
    auto acquire = [&]() {
      if (store_pending) {
        if (wg_leader) bulk_wait_read0();
        named_bar_sync(bar_id, 128);
        store_pending = false;
      }
    };

    for (int64_t p = blockIdx.x; p < n_pairs; p += gridDim.x) {
      const int64_t tile = 2 * p + t;
      if (tile >= n_tiles) break;
      uint8_t *my_acts = acts ? acts + tile * (int64_t)kActsPerTile * CN_ACT_TILE_BYTES : nullptr;

      if (kFusion) {
        // ---- mean over views -> fp32 residual stream in TMEM, relu(mean) -> first MMA operand
        const int64_t b = tile / tiles_per_view, tt = tile % tiles_per_view;
        const float inv_nv = 1.f / (float)nv;
        acquire();
#pragma unroll 1
        for (int c4 = 0; c4 < 4; ++c4) {
          uint32_t r[32];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            for (int v = 0; v < nv; ++v) {
              const uint8_t *src = in_tiles + ((b * nv + v) * tiles_per_view + tt) * (int64_t)CN_ACT_TILE_BYTES +
                                   tile_chunk_off(row, c4 * 4 + q);
              const uint4 x = ldg_nc_u4(src);
              acc[0] += bf16_lo(x.x); acc[1] += bf16_hi(x.x); acc[2] += bf16_lo(x.y); acc[3] += bf16_hi(x.y);
              acc[4] += bf16_lo(x.z); acc[5] += bf16_hi(x.z); acc[6] += bf16_lo(x.w); acc[7] += bf16_hi(x.w);
            }
            float m[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              m[i] = acc[i] * inv_nv;
              r[q * 8 + i] = __float_as_uint(m[i]);
              m[i] = fmaxf(m[i], 0.f);
            }
            *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, c4 * 4 + q)) =
                make_uint4(pack_bf16x2(m[0], m[1]), pack_bf16x2(m[2], m[3]), pack_bf16x2(m[4], m[5]),
                           pack_bf16x2(m[6], m[7]));
          }
          tmem_st32(t_h + c4 * 32, r);
        }
        tmem_wait_st();
        publish(my_acts, true);
      } else {
        mbar_wait(smem_u32(&bars.h_ready[t]), n_h & 1u);
        ++n_h;
        tc_fence_after();
      }

      for (int blk = 0; blk < 3; ++blk) {
        if (!(kFusion && blk == 0)) {
          // E_a: relu(h + cumulative bias) -> operand of the block's first Dense
          acquire();
          epi_tmem_to_act<true>(t_h, tab + (2 * blk) * 128, act, row);
          publish(my_acts ? my_acts + (2 * blk) * CN_ACT_TILE_BYTES : nullptr, true);
        }
        // E_b: relu(t + ba) -> operand of the block's second Dense
        mbar_wait(smem_u32(&bars.tmp_ready[t]), n_tmp & 1u);
        ++n_tmp;
        tc_fence_after();
        acquire();
        epi_tmem_to_act<true>(t_tmp, tab + (2 * blk + 1) * 128, act, row);
        publish(my_acts ? my_acts + (2 * blk + 1) * CN_ACT_TILE_BYTES : nullptr, true);
        mbar_wait(smem_u32(&bars.h_ready[t]), n_h & 1u);
        ++n_h;
        tc_fence_after();
      }

      acquire();
      if (!kFusion) {
        // ---- h3 = h + cumulative bias -> bf16 tile for the fusion stage
        epi_tmem_to_act<false>(t_h, tab + 6 * 128, act, row);
        tc_fence_before();
        mbar_arrive(smem_u32(&bars.h_free[t]));
        publish(out_tiles + tile * (int64_t)CN_ACT_TILE_BYTES, false);
      } else {
        // ---- readout: relu(e) . Wr + br, sigmoid / softplus (layers.py:392-397)
        const float *wr = tab + 7 * 128;
        float y0 = wr[512 + 0], y1 = wr[512 + 1], y2 = wr[512 + 2], y3 = wr[512 + 3];
#pragma unroll 1
        for (int c4 = 0; c4 < 4; ++c4) {
          uint32_t r[32];
          tmem_ld32(t_h + c4 * 32, r);
          tmem_wait_ld();
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int col = c4 * 32 + q * 8 + i;
              v[i] = fmaxf(__uint_as_float(r[q * 8 + i]) + tab[6 * 128 + col], 0.f);
              const float4 w = *reinterpret_cast<const float4 *>(wr + col * 4);
              y0 += v[i] * w.x; y1 += v[i] * w.y; y2 += v[i] * w.z; y3 += v[i] * w.w;
            }
            if (my_acts)
              *reinterpret_cast<uint4 *>(act + tile_chunk_off(row, c4 * 4 + q)) =
                  make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                             pack_bf16x2(v[6], v[7]));
          }
        }
        float4 o;
        o.x = 1.f / (1.f + expf(-y0));
        o.y = 1.f / (1.f + expf(-y1));
        o.z = 1.f / (1.f + expf(-y2));
        o.w = y3 > 20.f ? y3 : log1pf(expf(y3));
        reinterpret_cast<float4 *>(rgbs)[tile * 128 + row] = o;
        publish(my_acts ? my_acts + 6 * CN_ACT_TILE_BYTES : nullptr, false);
      }
    }
    if (wg_leader) bulk_wait0();
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

static int launch_fwd(bool fusion, const void *packed, const void *in_tiles, int64_t n_tiles, int nv,
                      int64_t tiles_per_view, void *out_tiles, float *rgbs, void *acts, cudaStream_t st) {
  const int64_t n_pairs = (n_tiles + 1) / 2;
  const unsigned grid = (unsigned)(n_pairs < kNumSMs ? n_pairs : kNumSMs);
  cudaError_t e;
  if (fusion) {
    e = cudaFuncSetAttribute(mlp_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFwdSmemBytes);
    if (e == cudaSuccess)
      mlp_fwd_kernel<true><<<grid, kFwdThreads, kFwdSmemBytes, st>>>(
          (const uint8_t *)packed, (const uint8_t *)in_tiles, n_tiles, nv, tiles_per_view, (uint8_t *)out_tiles, rgbs,
          (uint8_t *)acts);
  } else {
    e = cudaFuncSetAttribute(mlp_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFwdSmemBytes);
    if (e == cudaSuccess)
      mlp_fwd_kernel<false><<<grid, kFwdThreads, kFwdSmemBytes, st>>>(
          (const uint8_t *)packed, (const uint8_t *)in_tiles, n_tiles, nv, tiles_per_view, (uint8_t *)out_tiles, rgbs,
          (uint8_t *)acts);
  }
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
  return launch_fwd(false, packed, x_tiles, n_tiles, 1, n_tiles, h3_tiles, nullptr, acts, (cudaStream_t)stream);
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
                    (cudaStream_t)stream);
}
