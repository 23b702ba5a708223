// Backward of the first Dense (379 -> 128, layers.py:363) on tcgen05 / TMEM: the two role bodies shared by the
// stand-alone kernels (cn_mlp_l0_dgrad / cn_mlp_l0_wgrad, mlp_l0.cu) and the chained stage kernel (mlp_stage_bwd.cu):
//   dgrad : dX = g0 . W0^T   per 128-row tile, written as a bf16 MLP-input tile image
//           (same column order as X) for cn_project_gather_bwd
//   wgrad : dW0 += X^T . g0  over all tiles; the three 128 x 128 accumulators (384 input columns) stay in TMEM
//           for the whole kernel; db0 = column sums of g0 by the otherwise idle epilogue warps, straight from the
//           shared-memory operand
// g0 arrives from a global buffer or, chained, through a TileLink from the role that ran block 0's backward.
#pragma once
#include "common.cuh"
#include "mlp_bwd_block.cuh"
#include "mlp_layout.cuh"

namespace cn {

// ============================================================================================
// dgrad
// ============================================================================================
constexpr int kDgThreads = 64 + 512;  // producer, MMA, 2 epilogue groups of 8 warps (4 lane quadrants x 2 column halves)
constexpr uint32_t kDgSmemW = 0;                          // 3 x 32 KiB, resident
constexpr uint32_t kDgSmemG = 3 * CN_ACT_TILE_BYTES;      // 2 x 32 KiB (double buffer)
constexpr uint32_t kDgSmemOut = 5 * CN_ACT_TILE_BYTES;    // 2 x 32 KiB staging for the bulk stores
constexpr uint32_t kDgSmemBytes = 7 * CN_ACT_TILE_BYTES + 1024;

struct DgBars {
  uint64_t w_full;
  uint64_t g_full[2], g_free[2];
  uint64_t acc_ready[4], acc_free[4];
};

struct L0DgradArgs {
  const uint8_t *w0d;        // dgrad images of W0 (3 x 32 KiB)
  const uint8_t *g0_tiles;   // g0 from a global buffer (in.ring == nullptr)
  int64_t n_tiles;
  uint8_t *dx_tiles;
  float scale;
  TileLink in;
};

__device__ __forceinline__ void l0_dgrad_init_barriers(DgBars &bars) {
  mbar_init(smem_u32(&bars.w_full), 1);
  for (int i = 0; i < 2; ++i) {
    mbar_init(smem_u32(&bars.g_full[i]), 1);
    mbar_init(smem_u32(&bars.g_free[i]), 1);
  }
  for (int i = 0; i < 4; ++i) {
    mbar_init(smem_u32(&bars.acc_ready[i]), 1);
    mbar_init(smem_u32(&bars.acc_free[i]), 8);
  }
}

// role body: warps 0 .. kDgThreads/32 - 1 of the CTA (later warps fall through); barriers initialised, 512 TMEM
// columns at `tmem`, CTA synchronised
__device__ __forceinline__ void l0_dgrad_role(const L0DgradArgs &A, DgBars &bars, uint8_t *smem, uint32_t tmem, int cta,
                                              int n_cta) {
  const uint32_t s_base = smem_u32(smem);
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int64_t n_tiles = A.n_tiles;
  const float scale = A.scale;
  uint8_t *__restrict__ dx_tiles = A.dx_tiles;
  if (warp >= kDgThreads / 32) return;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(smem_u32(&bars.w_full), 3 * CN_ACT_TILE_BYTES);
      for (int c = 0; c < 3; ++c)
        bulk_g2s(s_base + kDgSmemW + c * CN_ACT_TILE_BYTES, A.w0d + c * CN_ACT_TILE_BYTES, CN_ACT_TILE_BYTES,
                 smem_u32(&bars.w_full));
      uint32_t it = 0;
      for (int64_t tile = cta; tile < n_tiles; tile += n_cta, ++it) {
        const uint32_t buf = it & 1u;
        mbar_wait(smem_u32(&bars.g_free[buf]), ((it >> 1) & 1u) ^ 1u);
        const uint8_t *src = A.in.ring ? link_wait_filled(A.in, tile) : A.g0_tiles + tile * (int64_t)CN_ACT_TILE_BYTES;
        mbar_expect_tx(smem_u32(&bars.g_full[buf]), CN_ACT_TILE_BYTES);
        bulk_g2s(s_base + kDgSmemG + buf * CN_ACT_TILE_BYTES, src, CN_ACT_TILE_BYTES, smem_u32(&bars.g_full[buf]));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(128, 128, 0, 0);
      mbar_wait(smem_u32(&bars.w_full), 0);
      uint32_t it = 0, unit = 0;
      for (int64_t tile = cta; tile < n_tiles; tile += n_cta, ++it) {
        const uint32_t buf = it & 1u;
        mbar_wait(smem_u32(&bars.g_full[buf]), (it >> 1) & 1u);
        if (A.in.ring) link_release(A.in, tile);              // g0 is in shared memory: its ring slot is free
        for (int c = 0; c < 3; ++c, ++unit) {
          const uint32_t slot = unit & 3u;
          mbar_wait(smem_u32(&bars.acc_free[slot]), ((unit >> 2) & 1u) ^ 1u);
          tc_fence_after();
#pragma unroll
          for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int j = 0; j < 4; ++j)
              umma_f16(tmem + slot * 128,
                       umma_desc(s_base + kDgSmemG + buf * CN_ACT_TILE_BYTES + s * CN_SLAB_BYTES + j * kKmajKstep,
                                 kKmajLbo, kSboBytes),
                       umma_desc(s_base + kDgSmemW + c * CN_ACT_TILE_BYTES + s * CN_SLAB_BYTES + j * kKmajKstep,
                                 kKmajLbo, kSboBytes),
                       idesc, (s | j) ? 1u : 0u);
          umma_commit(smem_u32(&bars.acc_ready[slot]));
        }
        umma_commit(smem_u32(&bars.g_free[buf]));
      }
    }
  } else {
    // Two epilogue groups of 8 warps take the accumulators alternately (unit u -> group u % 2, staging buffer u % 2):
    // a single epilogue is a latency chain (tcgen05.ld -> convert -> st.shared -> barrier -> bulk store, ~1400 cycles
    // with 8 warps) three times longer than the 8 MMAs that feed it, two of them in flight hide half of it
    const uint32_t quad = warp & 3u;
    const uint32_t grp = (warp - 2) >> 3;                     // epilogue group 0 / 1
    const uint32_t half = ((warp - 2) >> 2) & 1u;
    const uint32_t row = quad * 32u + lane;
    const uint32_t lane_off = (quad * 32u) << 16;
    const bool leader = (((warp - 2) & 7u) == 0 && lane == 0);
    uint8_t *out = smem + kDgSmemOut + grp * CN_ACT_TILE_BYTES;
    uint32_t unit = 0;
    for (int64_t tile = cta; tile < n_tiles; tile += n_cta) {
      for (int c = 0; c < 3; ++c, ++unit) {
        if ((unit & 1u) != grp) continue;
        const uint32_t slot = unit & 3u;
        mbar_wait(smem_u32(&bars.acc_ready[slot]), (unit >> 2) & 1u);
        tc_fence_after();
        uint32_t r[64];
        tmem_ld32(tmem + lane_off + slot * 128 + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
        tmem_ld32(tmem + lane_off + slot * 128 + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
        // this group's previous store (two units ago) has been read out of the staging buffer
        if (leader) bulk_wait_read0();
        named_bar_sync(1 + grp, 256);
        tmem_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars.acc_free[slot]));
#pragma unroll
        for (int q = 0; q < 8; ++q)
          *reinterpret_cast<uint4 *>(out + tile_chunk_off(row, half * 8 + q)) = make_uint4(
              cvt_bf16x2(__uint_as_float(r[q * 8 + 0]) * scale, __uint_as_float(r[q * 8 + 1]) * scale),
              cvt_bf16x2(__uint_as_float(r[q * 8 + 2]) * scale, __uint_as_float(r[q * 8 + 3]) * scale),
              cvt_bf16x2(__uint_as_float(r[q * 8 + 4]) * scale, __uint_as_float(r[q * 8 + 5]) * scale),
              cvt_bf16x2(__uint_as_float(r[q * 8 + 6]) * scale, __uint_as_float(r[q * 8 + 7]) * scale));
        fence_proxy_async_smem();
        named_bar_sync(1 + grp, 256);
        if (leader) {
          bulk_s2g(dx_tiles + tile * (int64_t)CN_X_TILE_BYTES + c * CN_ACT_TILE_BYTES,
                   s_base + kDgSmemOut + grp * CN_ACT_TILE_BYTES, CN_ACT_TILE_BYTES);
          bulk_commit();
        }
      }
    }
    if (leader) bulk_wait0();
  }
}

// ============================================================================================
// wgrad
// ============================================================================================
constexpr int kWgThreads = 192;
constexpr int kWgSlots = 14;  // 16 KiB slabs in pairs; a tile = 1 (g0) + 3 (X) pairs: 1.75 tiles in flight
constexpr uint32_t kWgSmemBytes = kWgSlots * CN_SLAB_BYTES + 1024;

struct WgBars {
  uint64_t full[kWgSlots / 2], empty[kWgSlots / 2];  // one barrier per slab PAIR
  uint64_t all_done;
};

struct L0WgradArgs {
  const uint8_t *x_tiles;
  const uint8_t *g0_tiles;   // g0 from a global buffer (in.ring == nullptr)
  int64_t n_tiles;
  float *g_w0, *g_b0;
  float scale;
  TileLink in;
};

__device__ __forceinline__ void l0_wgrad_init_barriers(WgBars &bars) {
  for (int i = 0; i < kWgSlots / 2; ++i) {
    mbar_init(smem_u32(&bars.full[i]), 1);
    mbar_init(smem_u32(&bars.empty[i]), 1 + 128);  // tcgen05.commit + the 128 bias-sum threads
  }
  mbar_init(smem_u32(&bars.all_done), 1);
}

// role body: warps 0 .. kWgThreads/32 - 1 of the CTA (later warps fall through)
__device__ __forceinline__ void l0_wgrad_role(const L0WgradArgs &A, WgBars &bars, uint8_t *smem, uint32_t tmem, int cta,
                                              int n_cta) {
  const uint32_t s_base = smem_u32(smem);
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  constexpr int kPairs = kWgSlots / 2;
  const int64_t n_tiles = A.n_tiles;
  const float scale = A.scale;
  const uint8_t *__restrict__ x_tiles = A.x_tiles;
  float *__restrict__ g_w0 = A.g_w0, *__restrict__ g_b0 = A.g_b0;
  if (warp >= kWgThreads / 32) return;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t pc = 0;  // pair counter
      auto load_pair = [&](const uint8_t *src) {
        const uint32_t ps = pc % kPairs;
        mbar_wait(smem_u32(&bars.empty[ps]), ((pc / kPairs) & 1u) ^ 1u);
        mbar_expect_tx(smem_u32(&bars.full[ps]), CN_ACT_TILE_BYTES);
        bulk_g2s(s_base + ps * CN_ACT_TILE_BYTES, src, CN_ACT_TILE_BYTES, smem_u32(&bars.full[ps]));
        ++pc;
      };
      for (int64_t tile = cta; tile < n_tiles; tile += n_cta) {
        load_pair(A.in.ring ? link_wait_filled(A.in, tile) : A.g0_tiles + tile * (int64_t)CN_ACT_TILE_BYTES);
        for (int c = 0; c < 3; ++c) load_pair(x_tiles + tile * (int64_t)CN_X_TILE_BYTES + c * CN_ACT_TILE_BYTES);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_mn = umma_idesc(128, 128, 1, 1);
      uint32_t cc = 0;
      bool first = true;
      auto next_pair = [&]() -> uint32_t {
        const uint32_t ps = cc % kPairs;
        mbar_wait(smem_u32(&bars.full[ps]), (cc / kPairs) & 1u);
        ++cc;
        return ps;
      };
      for (int64_t tile = cta; tile < n_tiles; tile += n_cta) {
        const uint32_t pg = next_pair();
        if (A.in.ring) link_release(A.in, tile);              // g0 is in shared memory: its ring slot is free
        for (int c = 0; c < 3; ++c) {
          const uint32_t px = next_pair();                    // the chunk's MMAs start as soon as ITS X pair has landed
          tc_fence_after();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            umma_f16(tmem + c * 128, umma_desc(s_base + px * CN_ACT_TILE_BYTES + j * kMNmajKstep, kMNmajLbo, kSboBytes),
                     umma_desc(s_base + pg * CN_ACT_TILE_BYTES + j * kMNmajKstep, kMNmajLbo, kSboBytes), idesc_mn,
                     (!first || j) ? 1u : 0u);
          umma_commit(smem_u32(&bars.empty[px]));
        }
        umma_commit(smem_u32(&bars.empty[pg]));
        first = false;
      }
      umma_commit(smem_u32(&bars.all_done));
    }
  } else {
    const uint32_t quad = warp & 3u;
    const uint32_t row = quad * 32u + lane;
    const uint32_t lane_off = (quad * 32u) << 16;
    // ---- db0 = column sums of g0, read from the shared-memory operand while the tensor pipe works:
    // warp q sums rows [32q, 32q+32), lane l the columns 2l, 2l+1 of both slabs
    {
      float b00 = 0.f, b01 = 0.f, b10 = 0.f, b11 = 0.f;
      const uint32_t chunk = lane >> 2, within = (lane & 3u) * 4u;
      uint32_t cc = 0;
      for (int64_t tile = cta; tile < n_tiles; tile += n_cta) {
        for (int p = 0; p < 4; ++p, ++cc) {
          const uint32_t ps = cc % kPairs;
          mbar_wait(smem_u32(&bars.full[ps]), (cc / kPairs) & 1u);
          if (p == 0) {                                  // the g0 pair comes first in a tile's four pairs
            const uint8_t *g = smem + ps * CN_ACT_TILE_BYTES + within;
#pragma unroll 8
            for (uint32_t r = quad * 32u; r < quad * 32u + 32u; ++r) {
              const uint32_t off = r * 128u + ((chunk ^ (r & 7u)) << 4);
              const uint32_t v0 = *reinterpret_cast<const uint32_t *>(g + off);
              const uint32_t v1 = *reinterpret_cast<const uint32_t *>(g + CN_SLAB_BYTES + off);
              b00 += bf16_lo(v0); b01 += bf16_hi(v0);
              b10 += bf16_lo(v1); b11 += bf16_hi(v1);
            }
          }
          mbar_arrive(smem_u32(&bars.empty[ps]));
        }
      }
      atomicAdd(&g_b0[2 * lane], b00 * scale);
      atomicAdd(&g_b0[2 * lane + 1], b01 * scale);
      atomicAdd(&g_b0[64 + 2 * lane], b10 * scale);
      atomicAdd(&g_b0[64 + 2 * lane + 1], b11 * scale);
    }
    mbar_wait(smem_u32(&bars.all_done), 0);
    tc_fence_after();
#pragma unroll 1
    for (int c = 0; c < 3; ++c) {
      const int ref = x_col_to_ref(c * 128 + row);  // accumulator row = internal input column
#pragma unroll 1
      for (int c4 = 0; c4 < 4; ++c4) {
        uint32_t r[32];
        tmem_ld32(tmem + lane_off + c * 128 + c4 * 32, r);
        tmem_wait_ld();
        if (ref >= 0) {
          float *dst = g_w0 + ref * 128 + c4 * 32;
#pragma unroll
          for (int q = 0; q < 8; ++q)
            red_add_f4(dst + q * 4, __uint_as_float(r[q * 4 + 0]) * scale, __uint_as_float(r[q * 4 + 1]) * scale,
                       __uint_as_float(r[q * 4 + 2]) * scale, __uint_as_float(r[q * 4 + 3]) * scale);
        }
      }
    }
  }
}

}  // namespace cn
