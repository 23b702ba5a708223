// cn_mlp_stage_bwd: the backward of one MLP STAGE -- three ResNet blocks and, for the per-view stage, the first
// Dense's dgrad + wgrad -- as ONE launch of role-specialised persistent CTAs chained through L2-resident tile rings.
//
// Why.  Launched block by block (cn_mlp_block_bwd x 3, cn_mlp_l0_wgrad, cn_mlp_l0_dgrad) the gradient stream g makes a
// round trip through HBM between every two kernels: a block reads A1 + g_y and writes g_h (96 KiB per tile), the
// first Dense reads g0 twice.  The kernels are bound by exactly that traffic (59 ... 94 % of the measured HBM peak in
// the ncu captures of round 1).  The weights of three blocks (192 KiB) and their six weight-gradient accumulators
// (768 TMEM columns) do not fit one SM, so the stage is not fused into one CTA but PIPELINED ACROSS CTAs:
//
//   role 0: block k+2   reads g_y (global)  -> ring 0
//   role 1: block k+1   ring 0              -> ring 1
//   role 2: block k     ring 1              -> ring 2 (or g_h global)
//   role 3: first Dense dgrad   ring 2 -> dX tiles (global)            } per-view stage only;
//   role 4: first Dense wgrad   ring 2 (+ X tiles, global) -> dW0, db0 } both read every tile of ring 2
//
// Every CTA is one of the existing role bodies (mlp_bwd_block.cuh, mlp_l0.cuh) with its own resident weights and
// TMEM-resident weight gradients; role r owns n_r CTAs and tile t is worked on by CTA t % n_r of every role.  A ring is
// kRingSlots tile slots per producer CTA in global memory (12 ... 19 MiB in total: it lives in the 126 MB L2, the g
// stream never reaches HBM) with two counters per slot:
//   filled   += 1  by the producer once its bulk store has completed (cp.async.bulk.wait_group 0, proxy fence,
//                  __threadfence, red.release.gpu)
//   released += 1  by every consumer once its bulk load has landed in shared memory (red.release.gpu)
// The producer of tile t waits for `released` of the slot's previous tile, the consumer for `filled` of tile t
// (ld.acquire.gpu spins of one thread, bounded by the watchdog).  All waits point at strictly older tiles of the same
// launch or at the producer of the same tile, so the schedule cannot deadlock as long as every CTA is resident: the
// grid is at most one CTA per SM.
// HBM traffic per tile of a stage: 3 x A1 + g_y + g_h = 160 KiB instead of 288 KiB (fusion stage); per-view stage with
// the first Dense: 3 x A1 + g_y(shared by the views) + X + dX instead of that plus five g round trips.
#include <cstdlib>

#include "mlp_bwd_block.cuh"
#include "mlp_l0.cuh"

namespace cn {

struct StageArgs {
  BlockBwdArgs blk[3];      // in backward order: blk[0] is the stage's last block
  L0DgradArgs dg;
  L0WgradArgs wg;
  int n_blk_cta;            // CTAs per block role
  int n_dg_cta, n_wg_cta;   // 0 = role absent
  long long *stats;         // nullable: [grid][4] (role, cycles, cycles waiting upstream, cycles waiting downstream)
};

union StageBars {
  BwdBars blk;
  DgBars dg;
  WgBars wg;
};

__global__ void __launch_bounds__(kBwdThreads, 1) mlp_stage_bwd_kernel(const __grid_constant__ StageArgs S) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) StageBars bars;
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float s_ba[128];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t warp = threadIdx.x >> 5;
  const int bid = (int)blockIdx.x;
  const int n_blk_total = 3 * S.n_blk_cta;
  // role of this CTA
  int role, cta;
  if (bid < n_blk_total) {
    role = bid / S.n_blk_cta;
    cta = bid % S.n_blk_cta;
  } else if (bid < n_blk_total + S.n_dg_cta) {
    role = 3;
    cta = bid - n_blk_total;
  } else {
    role = 4;
    cta = bid - n_blk_total - S.n_dg_cta;
  }
  if (role < 3 && threadIdx.x < 128) s_ba[threadIdx.x] = S.blk[role].bias_a[threadIdx.x];
  if (threadIdx.x == 0) {
    if (role < 3) block_bwd_init_barriers(bars.blk);
    else if (role == 3) l0_dgrad_init_barriers(bars.dg);
    else l0_wgrad_init_barriers(bars.wg);
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
  const long long c0 = clock64();
  if (role < 3) {
    BlockBwdArgs A = S.blk[role];
    A.stats = S.stats ? S.stats + (size_t)bid * 4 + 1 : nullptr;
    block_bwd_role(A, bars.blk, s_ba, smem, tmem, cta, S.n_blk_cta);
  } else if (role == 3) {
    l0_dgrad_role(S.dg, bars.dg, smem, tmem, cta, S.n_dg_cta);
  } else {
    l0_wgrad_role(S.wg, bars.wg, smem, tmem, cta, S.n_wg_cta);
  }
  if (S.stats && threadIdx.x == 0) {
    S.stats[(size_t)bid * 4] = role;
    if (role >= 3) S.stats[(size_t)bid * 4 + 1] = clock64() - c0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---- role split and workspace layout (host)
struct StagePlan {
  int n_blk, n_dg, n_wg;
  size_t counters_bytes, ring_bytes[3], total;
};

static StagePlan stage_plan(int64_t n_tiles, bool with_dgrad, bool with_wgrad) {
  StagePlan p{};
  const int sms = num_sms();
  // relative cost of a tile in a block role / the first Dense's dgrad / wgrad role (measured stand-alone, B200: 5
  // GEMMs + three epilogues vs 3 GEMMs + a 96 KiB store vs 3 GEMMs): CTAs are split in that proportion
  double cb = 5.5, cd = with_dgrad ? 3.0 : 0.0, cw = with_wgrad ? 2.6 : 0.0;
  if (const char *e = getenv("CN_STAGE_COST")) {    // tuning aid: "block,dgrad,wgrad"
    double a, b, c;
    if (sscanf(e, "%lf,%lf,%lf", &a, &b, &c) == 3) {
      cb = a;
      cd = with_dgrad ? b : 0.0;
      cw = with_wgrad ? c : 0.0;
    }
  }
  const double tot = 3 * cb + cd + cw;
  p.n_blk = (int)(sms * cb / tot);
  if (p.n_blk < 1) p.n_blk = 1;
  int rest = sms - 3 * p.n_blk;
  if (with_dgrad && with_wgrad) {
    p.n_dg = (int)(rest * cd / (cd + cw) + 0.5);
    if (p.n_dg < 1) p.n_dg = 1;
    p.n_wg = rest - p.n_dg;
    if (p.n_wg < 1) p.n_wg = 1;
  } else if (with_dgrad) {
    p.n_dg = rest > 0 ? rest : 1;
  } else if (with_wgrad) {
    p.n_wg = rest > 0 ? rest : 1;
  }
  if (p.n_blk > n_tiles) p.n_blk = (int)n_tiles;
  if (p.n_dg > n_tiles) p.n_dg = (int)n_tiles;
  if (p.n_wg > n_tiles) p.n_wg = (int)n_tiles;
  // counters: 3 links x n_blk producers x kRingSlots x {filled, released}
  p.counters_bytes = ((size_t)3 * p.n_blk * kRingSlots * 2 * sizeof(uint32_t) + 255) & ~(size_t)255;
  size_t o = p.counters_bytes;
  for (int l = 0; l < 3; ++l) {
    p.ring_bytes[l] = (size_t)p.n_blk * kRingSlots * CN_ACT_TILE_BYTES;
    o += p.ring_bytes[l];
  }
  p.total = o;
  return p;
}

// diagnostics (tools/trace_stage_bwd.py): device buffer [<= 160 CTAs][4] long long that the next launches fill
static long long *g_stage_stats = nullptr;

}  // namespace cn

extern "C" int cn_debug_stage_stats(long long *device_buffer) {
  cn::g_stage_stats = device_buffer;
  return 0;
}

extern "C" size_t cn_mlp_stage_bwd_ws_bytes(void) {
  // an upper bound independent of the tile count and of which roles run (the split only shrinks n_blk)
  return cn::stage_plan((int64_t)1 << 40, false, false).total;
}

extern "C" int cn_mlp_stage_bwd(const void *packed, int32_t stage, const void *acts, int32_t act_stride_tiles,
                                int64_t n_tiles, const void *gy_tiles, int32_t gy_views, int64_t gy_tiles_per_view,
                                void *gh_tiles, const void *x_tiles, void *dx_tiles, float scale, float *grad, void *ws,
                                size_t ws_bytes, void *stream) {
  using namespace cn;
  CN_REQUIRE(packed && acts && gy_tiles && grad && ws, "cn_mlp_stage_bwd: null pointer");
  CN_REQUIRE(stage == 0 || stage == 1, "cn_mlp_stage_bwd: stage %d (0 = per-view blocks 0-2, 1 = fusion blocks 3-5)", stage);
  CN_REQUIRE(act_stride_tiles >= 3, "cn_mlp_stage_bwd: act_stride_tiles=%d", act_stride_tiles);
  CN_REQUIRE(n_tiles >= 0, "cn_mlp_stage_bwd: n_tiles %lld", (long long)n_tiles);
  CN_REQUIRE(gy_views >= 1 && (gy_views == 1 || (gy_tiles_per_view > 0 && n_tiles % ((int64_t)gy_views * gy_tiles_per_view) == 0)),
             "cn_mlp_stage_bwd: gy_views=%d gy_tiles_per_view=%lld", gy_views, (long long)gy_tiles_per_view);
  const bool with_wgrad = x_tiles != nullptr, with_dgrad = dx_tiles != nullptr;
  CN_REQUIRE(stage == 0 || (!with_wgrad && !with_dgrad), "cn_mlp_stage_bwd: the first Dense belongs to the per-view stage");
  CN_REQUIRE((with_wgrad || with_dgrad) != (gh_tiles != nullptr),
             "cn_mlp_stage_bwd: the stage's output gradient goes EITHER to gh_tiles OR to the first Dense roles "
             "(x_tiles / dx_tiles)");
  CN_REQUIRE((((uintptr_t)packed | (uintptr_t)acts | (uintptr_t)gy_tiles | (uintptr_t)gh_tiles | (uintptr_t)x_tiles |
               (uintptr_t)dx_tiles) & 127) == 0 && ((uintptr_t)grad & 15) == 0 && ((uintptr_t)ws & 255) == 0,
             "cn_mlp_stage_bwd: alignment");
  if (n_tiles == 0) return 0;
  const StagePlan P = stage_plan(n_tiles, with_dgrad, with_wgrad);
  CN_REQUIRE(ws_bytes >= P.total, "cn_mlp_stage_bwd: workspace too small (%zu < %zu bytes)", ws_bytes, P.total);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(ws, 0, P.counters_bytes, st);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(mlp_stage_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBwdSmemBytes);
  if (e != cudaSuccess) {
    set_error("cn_mlp_stage_bwd: %s", cudaGetErrorString(e));
    return (int)e;
  }
  static_assert(kBwdSmemBytes >= kDgSmemBytes && kBwdSmemBytes >= kWgSmemBytes, "dynamic shared memory of the roles");
  static_assert(kBwdThreads >= kDgThreads && kBwdThreads >= kWgThreads, "threads of the roles");

  const uint8_t *pk = (const uint8_t *)packed;
  uint32_t *counters = (uint32_t *)ws;
  uint8_t *rings = (uint8_t *)ws + P.counters_bytes;
  TileLink link[3];
  for (int l = 0; l < 3; ++l) {
    link[l].ring = rings;
    rings += P.ring_bytes[l];
    link[l].filled = counters + (size_t)(2 * l) * P.n_blk * kRingSlots;
    link[l].released = counters + (size_t)(2 * l + 1) * P.n_blk * kRingSlots;
    link[l].n_prod = P.n_blk;
    link[l].n_cons = 1;
  }
  link[2].n_cons = (with_dgrad ? 1 : 0) + (with_wgrad ? 1 : 0);
  const TileLink none{nullptr, nullptr, nullptr, 1, 1};

  StageArgs S{};
  const int base = stage == 0 ? 0 : 3;
  for (int r = 0; r < 3; ++r) {
    const int blk = base + 2 - r;
    S.blk[r] = block_bwd_args(pk, blk, (const uint8_t *)acts, act_stride_tiles, 2 - r, n_tiles, scale, grad);
    S.blk[r].in = r == 0 ? none : link[r - 1];
    S.blk[r].out = (r < 2 || with_dgrad || with_wgrad) ? link[r] : none;
  }
  S.blk[0].gy_tiles = (const uint8_t *)gy_tiles;
  S.blk[0].gy_views = gy_views;
  S.blk[0].gy_tpv = gy_tiles_per_view;
  S.blk[2].gh_tiles = (uint8_t *)gh_tiles;
  S.n_blk_cta = P.n_blk;
  S.n_dg_cta = with_dgrad ? P.n_dg : 0;
  S.n_wg_cta = with_wgrad ? P.n_wg : 0;
  if (with_dgrad) {
    S.dg.w0d = pk + kPackBwdW0;
    S.dg.n_tiles = n_tiles;
    S.dg.dx_tiles = (uint8_t *)dx_tiles;
    S.dg.scale = scale;
    S.dg.in = link[2];
  }
  if (with_wgrad) {
    S.wg.x_tiles = (const uint8_t *)x_tiles;
    S.wg.n_tiles = n_tiles;
    S.wg.g_w0 = grad + kParamW0;
    S.wg.g_b0 = grad + kParamB0;
    S.wg.scale = scale;
    S.wg.in = link[2];
  }
  S.stats = g_stage_stats;
  const unsigned grid = (unsigned)(3 * S.n_blk_cta + S.n_dg_cta + S.n_wg_cta);
  mlp_stage_bwd_kernel<<<grid, kBwdThreads, kBwdSmemBytes, st>>>(S);
  return check_launch("cn_mlp_stage_bwd");
}
