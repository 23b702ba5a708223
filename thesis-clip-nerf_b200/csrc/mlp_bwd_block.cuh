// Backward of one pre-activation ResNet block (layers.py:262-298) for a stream of 128-row tiles, on tcgen05 / TMEM --
// the role body shared by the stand-alone kernel (cn_mlp_block_bwd, mlp_bwd.cu) and by the chained stage kernel
// (cn_mlp_stage_bwd, mlp_stage_bwd.cu).        y = h + relu(relu(h) Wa + ba) Wb + bb
//   saved by the forward: A1 = relu(h) only (bf16 tile image).  The hidden activation is RECOMPUTED
//   here -- one extra GEMM per tile buys a third of the kernel's HBM traffic:
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
// GEMMs of a tile already move 320 KiB through shared memory, so the role is bound by SHARED-MEMORY TRAFFIC before
// it is bound by the tensor pipe.  Hence:
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
//
// Tile links.  The g stream of a tile may arrive from / leave to another ROLE of the same launch instead of a global
// buffer: a TileLink is a ring of tile slots in global memory (small enough to stay in L2: the bytes never reach HBM)
// with one fill counter and one release counter per slot; see mlp_stage_bwd.cu for the protocol.
#pragma once
#include "common.cuh"
#include "mlp_layout.cuh"

namespace cn {

// ---------------------------------------------------------------- tile links between roles of one launch
constexpr int kRingSlots = 6;
struct TileLink {
  uint8_t *ring;       // [n_prod][kRingSlots][CN_ACT_TILE_BYTES]; nullptr = no link (plain global buffer)
  uint32_t *filled;    // [n_prod][kRingSlots] times the slot has been filled (release store by the producer)
  uint32_t *released;  // [n_prod][kRingSlots] times the slot has been released, summed over its consumers
  int n_prod;          // CTAs of the producing role: tile t is slot (t / n_prod) % kRingSlots of producer t % n_prod
  int n_cons;          // consumers per tile (2 for the first Dense's dgrad + wgrad roles)
};
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(uint32_t *p, uint32_t v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_relaxed_gpu_add(uint32_t *p, uint32_t v) {
  asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// optional per-CTA accounting of the time one thread spends waiting on tile links (tools/trace_stage_bwd.py)
struct LinkStats {
  long long wait_filled = 0, wait_released = 0;
};
// bounded spin (a protocol bug becomes a trap, not a hung GPU)
__device__ __forceinline__ void spin_until_ge(const uint32_t *p, uint32_t want, long long *acc = nullptr) {
  if (ld_acquire_gpu(p) >= want) return;
  const long long c0 = acc ? clock64() : 0;
  const uint64_t t0 = globaltimer_ns();
  while (ld_acquire_gpu(p) < want) {
    __nanosleep(64);
    if (globaltimer_ns() - t0 > CN_WATCHDOG_NS) {
      printf("clipnerf_b200: tile-link watchdog block %d thread %d want %u have %u\n", (int)blockIdx.x, (int)threadIdx.x,
             want, ld_acquire_gpu(p));
      __trap();
    }
  }
  if (acc) *acc += clock64() - c0;
}
// generic-proxy <-> async-proxy ordering for global memory touched by both bulk copies and flags
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
struct LinkSlot {
  int64_t idx;     // producer * kRingSlots + slot
  uint32_t use;    // how many times the slot had been used before this tile
};
__device__ __forceinline__ LinkSlot link_slot(const TileLink &l, int64_t tile) {
  const int64_t p = tile % l.n_prod, i = tile / l.n_prod;
  LinkSlot s;
  s.idx = p * kRingSlots + (i % kRingSlots);
  s.use = (uint32_t)(i / kRingSlots);
  return s;
}
// consumer side, before the bulk load of `tile`: wait until its producer has filled the slot
__device__ __forceinline__ const uint8_t *link_wait_filled(const TileLink &l, int64_t tile, LinkStats *st = nullptr) {
  const LinkSlot s = link_slot(l, tile);
  spin_until_ge(l.filled + s.idx, s.use + 1, st ? &st->wait_filled : nullptr);
  fence_proxy_async_all();
  return l.ring + s.idx * (int64_t)CN_ACT_TILE_BYTES;
}
// consumer side, once the bulk load of `tile` has landed in shared memory (observed through its mbarrier: the read
// of the slot is complete, nothing this thread wrote has to become visible -> a relaxed reduction, no fence)
__device__ __forceinline__ void link_release(const TileLink &l, int64_t tile) {
  red_relaxed_gpu_add(l.released + link_slot(l, tile).idx, 1u);
}
// producer side, before the bulk store of `tile`: wait until every consumer of the slot's previous tile has read it
__device__ __forceinline__ uint8_t *link_wait_released(const TileLink &l, int64_t tile, LinkStats *st = nullptr) {
  const LinkSlot s = link_slot(l, tile);
  spin_until_ge(l.released + s.idx, s.use * (uint32_t)l.n_cons, st ? &st->wait_released : nullptr);
  return l.ring + s.idx * (int64_t)CN_ACT_TILE_BYTES;
}
// producer side, after the bulk store of `tile` has COMPLETED (cp.async.bulk.wait_group 0)
__device__ __forceinline__ void link_publish(const TileLink &l, int64_t tile) {
  fence_proxy_async_all();
  __threadfence();
  red_release_gpu_add(l.filled + link_slot(l, tile).idx, 1u);
}

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
  uint64_t g2_ready;                   // commit behind G2 alone: E1 loads and masks g_a2 while WB is still reading A2
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

// arguments of one block-backward role
struct BlockBwdArgs {
  const uint8_t *w_a_img, *w_b_img;   // dgrad images of Wa, Wb (packed weights)
  const float *bias_a;
  const uint8_t *acts;                // saved relu(h) tile of tile 0 (tile stride act_stride_bytes)
  int64_t act_stride_bytes, n_tiles;
  const uint8_t *gy_tiles;            // g_y from a global buffer (in.ring == nullptr) ...
  int64_t gy_views, gy_tpv;           // ... with the mean over views folded in
  uint8_t *gh_tiles;                  // g_h to a global buffer (out.ring == nullptr)
  float scale;
  float *g_wa, *g_ba, *g_wb, *g_bb;
  TileLink in, out;                   // g_y from / g_h to another role of the launch
  long long *stats;                   // nullable: [3] cycles in the role, waiting for the upstream / downstream role
};

// role arguments of block `blk` (0..5) of a packed net; g_y source / g_h destination are left to the caller
inline BlockBwdArgs block_bwd_args(const uint8_t *packed, int blk, const uint8_t *acts, int act_stride_tiles,
                                   int act_index, int64_t n_tiles, float scale, float *grad) {
  BlockBwdArgs A{};
  A.w_a_img = packed + pack_bwd_block(blk, 0);
  A.w_b_img = packed + pack_bwd_block(blk, 1);
  A.bias_a = reinterpret_cast<const float *>(packed + kPackTables) + (blk < 3 ? kTabFeature : kTabFusion) +
             (2 * (blk % 3) + 1) * 128;
  A.acts = acts + (int64_t)act_index * CN_ACT_TILE_BYTES;
  A.act_stride_bytes = (int64_t)act_stride_tiles * CN_ACT_TILE_BYTES;
  A.n_tiles = n_tiles;
  A.gy_views = 1;
  A.scale = scale;
  A.g_wa = grad + param_block_w(blk, 0);
  A.g_ba = grad + param_block_b(blk, 0);
  A.g_wb = grad + param_block_w(blk, 1);
  A.g_bb = grad + param_block_b(blk, 1);
  return A;
}

__device__ __forceinline__ void block_bwd_init_barriers(BwdBars &bars) {
  mbar_init(smem_u32(&bars.w_full), 1);
  for (int b = 0; b < 2; ++b) {
    mbar_init(smem_u32(&bars.full_a1[b]), 1);
    mbar_init(smem_u32(&bars.full_gy[b]), 1);
    mbar_init(smem_u32(&bars.free_a1[b]), kBwdEpiWarps + 1);
    mbar_init(smem_u32(&bars.gy_staged[b]), kBwdEpiWarps);
  }
  mbar_init(smem_u32(&bars.t_ready), 1);
  mbar_init(smem_u32(&bars.g2wb_done), 1);
  mbar_init(smem_u32(&bars.g2_ready), 1);
  mbar_init(smem_u32(&bars.g1_ready), 1);
  mbar_init(smem_u32(&bars.a2_ready), kBwdEpiWarps);
  mbar_init(smem_u32(&bars.gt_ready), kBwdEpiWarps);
  mbar_init(smem_u32(&bars.a2buf_free), 1);
  mbar_init(smem_u32(&bars.acc1_free), kBwdEpiWarps);
  mbar_init(smem_u32(&bars.all_done), 1);
}

// The role body: CTA `cta` of `n_cta` CTAs of this role works on tiles cta, cta + n_cta, ...  Called by all
// kBwdThreads threads after the barriers are initialised, s_ba holds the 128 floats of bias_a, 512 TMEM columns are
// allocated at `tmem` and the CTA has synchronised.
// kWgrad = false: gradient w.r.t. the block input only (the pose-query backward of a frozen field,
// grasp_optimizer.py:158-184): no weight-gradient GEMMs, no bias column sums, no flush.
template <bool kWgrad = true>
__device__ __forceinline__ void block_bwd_role(const BlockBwdArgs &A, BwdBars &bars, const float *s_ba, uint8_t *smem,
                                               uint32_t tmem, int cta, int n_cta) {
  const uint32_t s_base = smem_u32(smem);
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const uint8_t *__restrict__ acts = A.acts;
  const int64_t act_stride_bytes = A.act_stride_bytes, n_tiles = A.n_tiles, gy_views = A.gy_views, gy_tpv = A.gy_tpv;
  const float scale = A.scale;
  const uint32_t t_acc0 = tmem, t_acc1 = tmem + 128, t_dwa = tmem + 256, t_dwb = tmem + 384;
  const int64_t n_mine = n_tiles > cta ? (n_tiles - cta + n_cta - 1) / n_cta : 0;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      mbar_expect_tx(smem_u32(&bars.w_full), 2 * CN_ACT_TILE_BYTES);
      bulk_g2s(s_base + kBwSmemWb, A.w_b_img, CN_ACT_TILE_BYTES, smem_u32(&bars.w_full));
      bulk_g2s(s_base + kBwSmemWa, A.w_a_img, CN_ACT_TILE_BYTES, smem_u32(&bars.w_full));
      auto load_a1 = [&](int64_t it) {
        const int64_t tile = cta + it * n_cta;
        const uint32_t b = (uint32_t)it & 1u;
        mbar_expect_tx(smem_u32(&bars.full_a1[b]), CN_ACT_TILE_BYTES);
        bulk_g2s(s_base + kBwSmemA1 + b * CN_ACT_TILE_BYTES, acts + tile * act_stride_bytes, CN_ACT_TILE_BYTES,
                 smem_u32(&bars.full_a1[b]));
      };
      LinkStats lst;
      const long long c_start = clock64();
      auto load_gy = [&](int64_t it) {
        const int64_t tile = cta + it * n_cta;
        const uint32_t b = (uint32_t)it & 1u;
        const uint8_t *src;
        if (A.in.ring) {
          src = link_wait_filled(A.in, tile, &lst);  // the upstream role has produced this tile (it stays in L2)
        } else {
          // mean over views folded in: every view tile of a (scene, row tile) reads the same fusion-stage g tile
          const int64_t st = gy_views > 1 ? (tile / (gy_views * gy_tpv)) * gy_tpv + tile % gy_tpv : tile;
          src = A.gy_tiles + st * (int64_t)CN_ACT_TILE_BYTES;
        }
        mbar_expect_tx(smem_u32(&bars.full_gy[b]), CN_ACT_TILE_BYTES);
        bulk_g2s(s_base + kBwSmemGy + b * CN_ACT_TILE_BYTES, src, CN_ACT_TILE_BYTES, smem_u32(&bars.full_gy[b]));
      };
      for (int64_t it = 0; it < 2 && it < n_mine; ++it) {
        load_a1(it);
        load_gy(it);
      }
      for (int64_t it = 0; it < n_mine; ++it) {
        const int64_t tile = cta + it * n_cta;
        const uint32_t b = (uint32_t)it & 1u, ph2 = ((uint32_t)it >> 1) & 1u;
        // E2 of this tile has replaced g_y by g_h in place: stream it out, then refill the buffer
        mbar_wait(smem_u32(&bars.gy_staged[b]), ph2);
        CN_TRACE_AT(it, 24);
        if (A.out.ring && it > 0) {
          bulk_wait0();                               // the previous tile's store has completed (a tile time ago):
          link_publish(A.out, tile - n_cta);          // hand it to the next role
        }
        uint8_t *dst = A.out.ring ? link_wait_released(A.out, tile, &lst) : A.gh_tiles + tile * (int64_t)CN_ACT_TILE_BYTES;
        bulk_s2g(dst, s_base + kBwSmemGy + b * CN_ACT_TILE_BYTES, CN_ACT_TILE_BYTES);
        bulk_commit();
        if (it + 2 < n_mine) {
          mbar_wait(smem_u32(&bars.free_a1[b]), ph2);
          CN_TRACE_AT(it, 26);
          load_a1(it + 2);
          bulk_wait_read0();                          // the store has read the buffer: refill it
          load_gy(it + 2);                            // (may wait for the upstream role: last)
        }
      }
      bulk_wait0();
      if (A.out.ring && n_mine > 0) link_publish(A.out, cta + (n_mine - 1) * (int64_t)n_cta);
      if (A.stats) {
        A.stats[0] = clock64() - c_start;
        A.stats[1] = lst.wait_filled;
        A.stats[2] = lst.wait_released;
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      constexpr uint32_t idesc_kk = umma_idesc(128, 128, 0, 0);  // dgrad: both operands K-major
      constexpr uint32_t idesc_kn = umma_idesc(128, 128, 0, 1);  // recompute: A K-major, B MN-major
      constexpr uint32_t idesc_mn = umma_idesc(128, 128, 1, 1);  // wgrad: both operands MN-major
      // descriptors: one base per operand layout, built once; an operand at byte offset `off` of the CTA's shared memory
      // is base + (off >> 4) on the encoded start address (no carry: shared memory < 256 KiB).  Rebuilding the
      // descriptor from the address for every MMA costs the issuing thread ~20 % of the tensor rate when the
      // epilogue warps share its scheduler (tools/ubench_mma_issue.cu).
      const uint64_t d_k = umma_desc(s_base, kKmajLbo, kSboBytes), d_mn = umma_desc(s_base, kMNmajLbo, kSboBytes);
      auto gemm_kk = [&](uint32_t d_tmem, uint32_t a_off, uint32_t b_off) {
        const uint64_t da = d_k + (a_off >> 4), db = d_k + (b_off >> 4);
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
          for (int j = 0; j < 4; ++j)
            umma_f16(d_tmem, da + ((s * CN_SLAB_BYTES + j * kKmajKstep) >> 4), db + ((s * CN_SLAB_BYTES + j * kKmajKstep) >> 4),
                     idesc_kk, (s | j) ? 1u : 0u);
      };
      // t = A1 Wa with Wa stored [in][out] (K rows, N contiguous): K step 16 = 16 rows of the weight image
      auto gemm_t = [&](uint32_t a_off) {
        const uint64_t da = d_k + (a_off >> 4), db = d_mn + (kBwSmemWa >> 4);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          umma_f16(t_acc0, da + (((j >> 2) * CN_SLAB_BYTES + (j & 3) * kKmajKstep) >> 4), db + ((j * kMNmajKstep) >> 4), idesc_kn,
                   j ? 1u : 0u);
      };
      auto wgrad = [&](uint32_t d_tmem, uint32_t a_off, uint32_t b_off, bool first) {
        const uint64_t da = d_mn + (a_off >> 4), db = d_mn + (b_off >> 4);
#pragma unroll
        for (int j = 0; j < 8; ++j)  // 8 steps of 16 rows
          umma_f16(d_tmem, da + ((j * kMNmajKstep) >> 4), db + ((j * kMNmajKstep) >> 4), idesc_mn, (!first || j) ? 1u : 0u);
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
        umma_commit(smem_u32(&bars.g2_ready));
        CN_TRACE_AT(it, 1);
        mbar_wait(smem_u32(&bars.a2_ready), par);                 // (also: E0 has drained acc0)
        tc_fence_after();
        CN_TRACE_AT(it, 2);
        if (kWgrad) wgrad(t_dwb, kBwSmemA2, gy, it == 0);         // WB: dWb += A2^T g_y
        umma_commit(smem_u32(&bars.g2wb_done));
        CN_TRACE_AT(it, 3);
        // T of the next tile goes out behind WB, into the ~1 000 cycles the tensor pipe idles while E1 runs -- if its
        // A1 has landed by then.  E2 releases the A1 buffer before it waits for G1 (below), so the buffer is free as soon
        // as WA has read it and the load of tile it+2 lands ~1 000 ... 1 500 cycles into tile it+1: in time.  A T that
        // misses the window is issued after WA (not before it: a late WA releases A1 late, and the next load is late
        // again -- tools/trace_block_bwd.py showed that state sustaining itself).  The test is repeated for a bounded
        // time only: a T that starts late in E1 would hold up G1, which the epilogue is waiting for.  Measured
        // (tools/bench_block_bwd.py, one box, per-view / fusion blocks of cfg2): no polling 292.9 / 114.7 us, 300 cycles
        // 289.8 / 112.7, 600 -2.9 / -3.7 %, 900 -3.6 / -3.7 %, 1 200 and "until gt_ready" slower than 900 again.  With E1
        // loading g_a2 before WB is done (below) the order is the same: 0 / 300 / 600 / 900 cycles = 297.0 / 283.5 /
        // 282.7 / 280.6 us (per-view blocks; until gt_ready = 280.6).
#ifndef CN_BWD_POLL_CYCLES
#define CN_BWD_POLL_CYCLES 900
#endif
        bool next_issued = false;
        auto try_next_t = [&]() {
          if (has_next && !next_issued && mbar_try_wait(smem_u32(&bars.full_a1[nb]), nph2)) {
            tc_fence_after();
            gemm_t(kBwSmemA1 + nb * CN_ACT_TILE_BYTES);
            umma_commit(smem_u32(&bars.t_ready));
            next_issued = true;
          }
        };
        try_next_t();
        if (CN_BWD_POLL_CYCLES > 0 && has_next && !next_issued) {
          const long long c0 = clock64();
          while (!next_issued && clock64() - c0 < CN_BWD_POLL_CYCLES) try_next_t();
        }
        CN_TRACE_AT(it, 4);
        mbar_wait(smem_u32(&bars.gt_ready), par);
        tc_fence_after();
        CN_TRACE_AT(it, 5);
        gemm_kk(t_acc1, kBwSmemA2, kBwSmemWa);                    // G1: g_a1 = g_t Wa^T
        umma_commit(smem_u32(&bars.g1_ready));
        CN_TRACE_AT(it, 6);
        if (kWgrad) wgrad(t_dwa, a1, kBwSmemA2, it == 0);         // WA: dWa += A1^T g_t
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
      if (A.in.ring && threadIdx.x == 64) {
        // g_y of this tile is in shared memory (it landed long before t_ready): its ring slot is free.  Done by an
        // epilogue thread: the MMA-issuing thread is the critical one
        mbar_wait(smem_u32(&bars.full_gy[b]), ((uint32_t)it >> 1) & 1u);
        link_release(A.in, cta + it * (int64_t)n_cta);
      }
      // ---- E0: A2 = relu(t + ba) -> bf16 operand tile
      if (threadIdx.x == 64) CN_TRACE_AT(it, 18);
      mbar_wait(smem_u32(&bars.t_ready), par);
      if (threadIdx.x == 64) CN_TRACE_AT(it, 19);
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
      // ---- E1: g_t = g_a2 * [A2 > 0], over A2; dba += colsum(g_t).  Only the STORES wait for WB (it reads A2): g_a2 is
      // loaded and masked as soon as G2 alone has completed, while the MMA thread is still issuing WB
      mbar_wait(smem_u32(&bars.g2_ready), par);
      tc_fence_after();
      if (threadIdx.x == 64) CN_TRACE_AT(it, 12);
      {
        uint32_t r[32];
        tmem_ld32(t_acc1 + lane_off + cq * 32, r);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 16; ++j)
          a2w[j] = mul_bf16x2(cvt_bf16x2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1])), gt0_bf16x2(a2w[j]));
      }
      mbar_wait(smem_u32(&bars.g2wb_done), par);      // WB has finished reading A2
      if (threadIdx.x == 64) CN_TRACE_AT(it, 13);
#pragma unroll
      for (int q = 0; q < 4; ++q)
        *reinterpret_cast<uint4 *>(a2buf + choff[q]) =
            make_uint4(a2w[q * 4 + 0], a2w[q * 4 + 1], a2w[q * 4 + 2], a2w[q * 4 + 3]);
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars.gt_ready));
      if (threadIdx.x == 64) CN_TRACE_AT(it, 14);
      if (kWgrad) {
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
      if (kWgrad) {
        uint32_t t16[16];                              // the butterfly consumes its input: work on a copy, while G1 runs
#pragma unroll
        for (int j = 0; j < 16; ++j) t16[j] = gw[j];
        dbb += colsum_bf16x2(t16, lane);
      }
      // the ReLU mask of E2 comes from A1: its chunks are taken into registers NOW, and the buffer is released as far as
      // the epilogue is concerned -- it is then free as soon as WA has read it, ~1 300 cycles before the end of E2, and the
      // A1 of tile it+2 lands in time for T to be issued while E1 of tile it+1 runs (tools/trace_block_bwd.py)
      uint4 a[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) a[q] = *reinterpret_cast<const uint4 *>(a1buf + choff[q]);
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars.free_a1[b]));
      mbar_wait(smem_u32(&bars.g1_ready), par);
      tc_fence_after();
      if (threadIdx.x == 64) CN_TRACE_AT(it, 15);
      {
        uint32_t r[32];
        tmem_ld32(t_acc1 + lane_off + cq * 32, r);
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
      __syncwarp();                                  // every lane has written g_h
      if (lane == 0) mbar_arrive(smem_u32(&bars.gy_staged[b]));
      if (threadIdx.x == 64) CN_TRACE_AT(it, 17);
    }
    // ---- flush the weight gradients held in TMEM (each thread: its input unit, 32 output units)
    mbar_wait(smem_u32(&bars.all_done), 0);
    tc_fence_after();
    if (kWgrad && n_mine > 0) {
#pragma unroll 1
      for (int mtx = 0; mtx < 2; ++mtx) {
        float *gdst = (mtx == 0 ? A.g_wa : A.g_wb) + row * 128 + cq * 32;  // accumulator row = input unit
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
        atomicAdd((which == 0 ? A.g_ba : A.g_bb) + col, v * scale);
      }
    }
  }
}

}  // namespace cn
