// The encoder's final UpSampling2D(2, 'bilinear') folded into the gather (CnScene.feat_up == 2; layers.py:589,659,
// legacy_layers.py:190, model_without.py:76): the kernels below read / differentiate the [H/2, W/2, 256] map the
// upsample would have consumed.  Included by gather_binned.cu after the workspace / record definitions.
//
// Per axis, with f0 / a the clamped floor and weight of interpolate_bilinear on the 2x axis and k = f0 >> 1:
//     (1 - a) U[f0] + a U[f0 + 1] = sum_j w_j x[clamp(k + j - 1, 0, size - 1)]
//     f0 even: w = (0.25 (1 - a), 0.75, 0.25 a)        f0 odd: w = (0, 0.75 - 0.5 a, 0.25 + 0.5 a)
// (oracle/nerf_oracle.py: lowres_stencil, proven equal to the two-stage form in tests/test_oracle.py).  Rows are keyed
// by the base texel (kx, ky) -- 4x more rows per cell than on the upsampled grid -- and the sort makes the rows of a
// cell consecutive, so a warp that walks a contiguous piece of the sorted order keeps the 3x3 texel window in
// registers and touches shared memory / L2 only when the cell changes:
//   tile_rows_lr_kernel     8x8-cell tiles = 10x10 texels (100 KiB) of the low-resolution map staged by a producer
//                           warp; forward: 72 FMAs per lane and row out of the register window; backward: the nine dot
//                           products <dX, texel> combined with the weight derivatives
//   feature_grad_lr_kernel  one warp per (tile, cell row): every dX row is read ONCE and accumulated into a sliding
//                           3x3x256 register window; a window column is added to d_features (red.global.add.v4.f32)
//                           when the sweep has moved past it
#pragma once

namespace cn {

constexpr int kTL = kT + 2;                                    // texels per tile edge (one halo texel each side)
constexpr int kTileBytesLr = kTL * kTL * kTexBytes;            // 102400
constexpr uint32_t kRowsSlotBytesLr = kTileBytesLr + kChunkLr * 16;
constexpr uint32_t kRowsSmemLr = 2 * kRowsSlotBytesLr + 1024;
constexpr int kLrWarpsFwd = 16, kLrWarpsBwd = 12;              // consumer warps (the backward holds two dX batches)
template <bool kBackward>
constexpr int lr_consumer_warps() { return kBackward ? kLrWarpsBwd : kLrWarpsFwd; }
template <bool kBackward>
constexpr int lr_threads() { return 32 * (1 + lr_consumer_warps<kBackward>()); }

struct TileDescLr {
  int start, count;
  int cx0, cy0;   // first cell of the tile
  int x_lo, y_lo; // first texel held in shared memory
};

// weights of the three low-resolution taps of one axis and their derivative w.r.t. the interpolation weight
__device__ __forceinline__ void lr_weights(int odd, float a, float (&w)[3]) {
  w[0] = odd ? 0.f : 0.25f * (1.f - a);
  w[1] = odd ? 0.75f - 0.5f * a : 0.75f;
  w[2] = odd ? 0.25f + 0.5f * a : 0.25f * a;
}
__device__ __forceinline__ void lr_dweights(int odd, float (&dw)[3]) {
  dw[0] = odd ? 0.f : -0.25f;
  dw[1] = odd ? -0.5f : 0.f;
  dw[2] = odd ? 0.5f : 0.25f;
}

struct Window {
  float4 t[3][3][2];   // [tap row][tap column][channels 4l..4l+3 | 128+4l..128+4l+3]
};
__device__ __forceinline__ void load_window(Window &w, const float *tile, const TileDescLr &d, int loc, int Hl, int Wl,
                                            uint32_t l) {
  const int kx = d.cx0 + (loc & 0xff), ky = d.cy0 + ((loc >> 8) & 0xff);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int yy = min(max(ky + i - 1, 0), Hl - 1) - d.y_lo;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int xx = min(max(kx + j - 1, 0), Wl - 1) - d.x_lo;
      const float *t = tile + (yy * kTL + xx) * CN_FEAT_C;
      w.t[i][j][0] = *reinterpret_cast<const float4 *>(t + 4 * l);
      w.t[i][j][1] = *reinterpret_cast<const float4 *>(t + 128 + 4 * l);
    }
  }
}

__device__ __forceinline__ float4 f4_scale(float s, float4 a) { return make_float4(s * a.x, s * a.y, s * a.z, s * a.w); }
__device__ __forceinline__ float4 f4_fma(float s, float4 a, float4 c) {
  return make_float4(fmaf(s, a.x, c.x), fmaf(s, a.y, c.y), fmaf(s, a.z, c.z), fmaf(s, a.w, c.w));
}
__device__ __forceinline__ float f4_dot_bf16(float4 t, uint2 g) {
  return fmaf(t.x, bf16_lo(g.x), fmaf(t.y, bf16_hi(g.x), fmaf(t.z, bf16_lo(g.y), t.w * bf16_hi(g.y))));
}

template <bool kBackward>
__global__ void __launch_bounds__(lr_threads<kBackward>(), 1)
tile_rows_lr_kernel(CnScene sc, Ws g, int *__restrict__ cursor, uint8_t *__restrict__ x_tiles) {
  constexpr int kWarps = lr_consumer_warps<kBackward>();
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[2], bar_empty[2];
  __shared__ TileDescLr sdesc[2];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t l = lane_id();
  const int Hl = sc.H / 2, Wl = sc.W / 2;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), kWarps);
    }
    fence_barrier_init();
  }
  __syncthreads();
  const int n_items = g.hdr[kHdrItems];

  if (warp == 0) {
    // ============ producer: 10x10-texel tile of the low-resolution map + the item's row records ============
    if (l == 0) {
      int it = atomicAdd(cursor, 1);
      int4 item = it < n_items ? g.items[it] : make_int4(0, 0, 0, 0);
      for (uint32_t k = 0;; ++k) {
        const uint32_t slot = k & 1u;
        mbar_wait(smem_u32(&bar_empty[slot]), ((k >> 1) & 1u) ^ 1u);
        if (it >= n_items) {
          sdesc[slot].count = 0;
          mbar_arrive(smem_u32(&bar_full[slot]));
          break;
        }
        const int tx = item.x % g.tiles_x;
        const int rest = item.x / g.tiles_x;
        const int ty = rest % g.tiles_y;
        const int bv = rest / g.tiles_y;
        TileDescLr d;
        d.start = item.y;
        d.count = item.z;
        d.cx0 = tx * kT;
        d.cy0 = ty * kT;
        d.x_lo = max(d.cx0 - 1, 0);
        d.y_lo = max(d.cy0 - 1, 0);
        const int nx = min(d.cx0 + kT, Wl - 1) - d.x_lo + 1, ny = min(d.cy0 + kT, Hl - 1) - d.y_lo + 1;
        sdesc[slot] = d;
        const uint32_t row_bytes = (uint32_t)nx * kTexBytes;
        const uint32_t bar = smem_u32(&bar_full[slot]);
        mbar_expect_tx(bar, row_bytes * (uint32_t)ny + (uint32_t)item.z * 16u);
        const float *src = sc.features + (((int64_t)bv * Hl + d.y_lo) * Wl + d.x_lo) * CN_FEAT_C;
        const uint32_t dst = smem_u32(smem) + slot * kRowsSlotBytesLr;
        bulk_g2s(dst + kTileBytesLr, g.rec_s + item.y, (uint32_t)item.z * 16u, bar);
        for (int yy = 0; yy < ny; ++yy)
          bulk_g2s(dst + yy * (kTL * kTexBytes), src + (int64_t)yy * Wl * CN_FEAT_C, row_bytes, bar);
        it = atomicAdd(cursor, 1);
        item = it < n_items ? g.items[it] : make_int4(0, 0, 0, 0);
      }
    }
    return;
  }

  // ======== consumers: warp cw walks rows [cw * per, cw * per + per) of the item (sorted by cell) ========
  const uint32_t cw = warp - 1;
  for (uint32_t k = 0;; ++k) {
    const uint32_t slot = k & 1u;
    mbar_wait(smem_u32(&bar_full[slot]), (k >> 1) & 1u);
    const TileDescLr d = sdesc[slot];
    if (d.count == 0) break;
    const float *tile = reinterpret_cast<const float *>(smem + slot * kRowsSlotBytesLr);
    const int4 *recs = reinterpret_cast<const int4 *>(smem + slot * kRowsSlotBytesLr + kTileBytesLr);
    const int per = (d.count + kWarps - 1) / kWarps;
    const int j0 = (int)cw * per, j1 = min(d.count, j0 + per);
    Window win;
    int cell = -1;
    if constexpr (!kBackward) {
      for (int j = j0; j < j1; ++j) {
        const int4 cur = recs[j];
        if ((cur.y & 0xffff) != cell) {
          cell = cur.y & 0xffff;
          load_window(win, tile, d, cell, Hl, Wl, l);
        }
        float wx[3], wy[3];
        lr_weights((cur.y >> 16) & 1, __int_as_float(cur.z), wx);
        lr_weights((cur.y >> 17) & 1, __int_as_float(cur.w), wy);
        float4 o0 = make_float4(0.f, 0.f, 0.f, 0.f), o1 = o0;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          float4 r0 = f4_scale(wx[0], win.t[i][0][0]), r1 = f4_scale(wx[0], win.t[i][0][1]);
          r0 = f4_fma(wx[1], win.t[i][1][0], r0);
          r1 = f4_fma(wx[1], win.t[i][1][1], r1);
          r0 = f4_fma(wx[2], win.t[i][2][0], r0);
          r1 = f4_fma(wx[2], win.t[i][2][1], r1);
          o0 = f4_fma(wy[i], r0, o0);
          o1 = f4_fma(wy[i], r1, o1);
        }
        uint8_t *xt = x_tiles + (int64_t)(cur.x >> 7) * CN_X_TILE_BYTES;
        const uint32_t rr = (uint32_t)cur.x & 127u;
        *reinterpret_cast<uint2 *>(xt + tile_chunk_off(rr, l >> 1) + (l & 1) * 8) =
            make_uint2(pack_bf16x2(o0.x, o0.y), pack_bf16x2(o0.z, o0.w));
        *reinterpret_cast<uint2 *>(xt + tile_chunk_off(rr, 16 + (l >> 1)) + (l & 1) * 8) =
            make_uint2(pack_bf16x2(o1.x, o1.y), pack_bf16x2(o1.z, o1.w));
      }
    } else {
      // batches of 4 rows; the 8 dX loads of batch n+1 are in flight during the dot products of batch n
      auto issue = [&](int jb, uint2 (&ga)[4], uint2 (&gb)[4]) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int4 r = recs[min(jb + u, j1 - 1)];
          const uint8_t *t = x_tiles + (int64_t)(r.x >> 7) * CN_X_TILE_BYTES;
          const uint32_t rr = (uint32_t)r.x & 127u;
          ga[u] = __ldg(reinterpret_cast<const uint2 *>(t + tile_chunk_off(rr, l >> 1) + (l & 1) * 8));
          gb[u] = __ldg(reinterpret_cast<const uint2 *>(t + tile_chunk_off(rr, 16 + (l >> 1)) + (l & 1) * 8));
        }
      };
      uint2 ga[4], gb[4];
      if (j0 < j1) issue(j0, ga, gb);
      for (int jb = j0; jb < j1; jb += 4) {
        uint2 na[4], nb[4];
        const bool more = jb + 4 < j1;
        if (more) issue(jb + 4, na, nb);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (jb + u >= j1) break;
          const int4 cur = recs[jb + u];
          if ((cur.y & 0xffff) != cell) {
            cell = cur.y & 0xffff;
            load_window(win, tile, d, cell, Hl, Wl, l);
          }
          const int ox = (cur.y >> 16) & 1, oy = (cur.y >> 17) & 1;
          float wx[3], wy[3], dwx[3], dwy[3];
          lr_weights(ox, __int_as_float(cur.z), wx);
          lr_weights(oy, __int_as_float(cur.w), wy);
          lr_dweights(ox, dwx);
          lr_dweights(oy, dwy);
          // value = sum_ij wy_i wx_j <dX, T_ij>:  d/d ax = sum_ij wy_i wx'_j s_ij,  d/d ay = sum_ij wy'_i wx_j s_ij
          float gax = 0.f, gay = 0.f;
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            float s[3];
#pragma unroll
            for (int j = 0; j < 3; ++j) s[j] = f4_dot_bf16(win.t[i][j][0], ga[u]) + f4_dot_bf16(win.t[i][j][1], gb[u]);
            gax = fmaf(wy[i], dwx[0] * s[0] + dwx[1] * s[1] + dwx[2] * s[2], gax);
            gay = fmaf(dwy[i], wx[0] * s[0] + wx[1] * s[1] + wx[2] * s[2], gay);
          }
          float v = (l < 16) ? gax : gay;
          v += __shfl_xor_sync(0xffffffffu, (l < 16) ? gay : gax, 16);
#pragma unroll
          for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          const float v_hi = __shfl_sync(0xffffffffu, v, 16);
          if (l == 0) g.gpix[cur.x] = make_float2(v, v_hi);
        }
        if (more) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            ga[u] = na[u];
            gb[u] = nb[u];
          }
        }
      }
    }
    __syncwarp();
    if (l == 0) mbar_arrive(smem_u32(&bar_empty[slot]));
  }
}

// ------------------------------------------------------------------------------------------
// feature-map gradient on the low-resolution map
// ------------------------------------------------------------------------------------------
constexpr int kLrMaxSets = 3;
constexpr int kLrThreads = 128;       // 4 warps, 3 CTAs per SM (<= 168 registers: the 72-float window + three record streams)
constexpr int kLrCtasPerSm = 3;
constexpr int kLrUnitCap = 1024;      // a (tile, cell row) with more rows is split ...
constexpr int kLrSplit = 512;         // ... into independent pieces of this many rows of one set

struct LrGradArgs {
  GradSet set[kLrMaxSets];
  int n_sets;
  int n_keys, tiles_x, tiles_y, n_units;
  int *cursor;
  int *heavy_n;       // [0] deferred pieces, [1] cursor of the second kernel
  int2 *heavy_list;   // (unit * 4 + set, first row of the piece)
  int heavy_cap;
};

// rows [r0, r1) of every set (sorted by cell; all inside cell row `ly` of tile `bin`) -> d_features
__device__ __forceinline__ void lr_sweep(const CnScene &sc, const LrGradArgs &a, int bin, int ly, const int (&r0)[kLrMaxSets],
                                         const int (&r1)[kLrMaxSets], float *__restrict__ d_features) {
  const uint32_t l = lane_id();
  const int Hl = sc.H / 2, Wl = sc.W / 2;
  const int tx = bin % a.tiles_x;
  const int rest = bin / a.tiles_x;
  const int ty = rest % a.tiles_y;
  const int bv = rest / a.tiles_y;
  const int ky = ty * kT + ly;
  float *rows_out[3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
    rows_out[i] = d_features + (((int64_t)bv * Hl + min(max(ky + i - 1, 0), Hl - 1)) * Wl) * CN_FEAT_C;

  float acc[3][3][8];   // [window column][tap row][channel]
  bool dirty[3] = {false, false, false};
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) acc[c][i][ch] = 0.f;
  int cur = 0;   // the window's middle column is cell `cur` of this cell row

  auto flush0 = [&]() {   // window column 0 -> texel column clamp(cell - 1)
    if (dirty[0]) {
      const int X = min(max(tx * kT + cur - 1, 0), Wl - 1);
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        float *o = rows_out[i] + (int64_t)X * CN_FEAT_C;
        red_add_f4(o + 4 * l, acc[0][i][0], acc[0][i][1], acc[0][i][2], acc[0][i][3]);
        red_add_f4(o + 128 + 4 * l, acc[0][i][4], acc[0][i][5], acc[0][i][6], acc[0][i][7]);
      }
    }
  };
  auto slide = [&]() {    // cell cur -> cur + 1
    flush0();
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        acc[0][i][ch] = acc[1][i][ch];
        acc[1][i][ch] = acc[2][i][ch];
        acc[2][i][ch] = 0.f;
      }
    dirty[0] = dirty[1];
    dirty[1] = dirty[2];
    dirty[2] = false;
    ++cur;
  };

  // one stream of records per set: lane f holds row base + f of the set's range
  const uint32_t lane_off = ((l >> 4) * CN_SLAB_BYTES) + (((l >> 1) & 7u) << 4) + (l & 1u) * 8u;
  int base[kLrMaxSets], n[kLrMaxSets], used[kLrMaxSets], cellx[kLrMaxSets];
  float wx0[kLrMaxSets], wx1[kLrMaxSets], wx2[kLrMaxSets], wy0[kLrMaxSets], wy1[kLrMaxSets], wy2[kLrMaxSets];
  uint32_t a_lo[kLrMaxSets], a_hi[kLrMaxSets];
  auto refill = [&](int s) {
    base[s] += n[s];
    n[s] = min(32, r1[s] - base[s]);
    used[s] = 0;
    cellx[s] = 99;
    if ((int)l < n[s]) {
      const int4 rec = __ldg(a.set[s].rec_s + base[s] + l);
      cellx[s] = rec.y & 0xff;
      float w[3];
      lr_weights((rec.y >> 16) & 1, __int_as_float(rec.z), w);
      wx0[s] = w[0]; wx1[s] = w[1]; wx2[s] = w[2];
      lr_weights((rec.y >> 17) & 1, __int_as_float(rec.w), w);
      wy0[s] = w[0]; wy1[s] = w[1]; wy2[s] = w[2];
      const uint32_t row = (uint32_t)rec.x;
      const uint64_t addr = (uint64_t)(uintptr_t)a.set[s].dx +
                            ((uint64_t)(row >> 7) * CN_X_TILE_BYTES + (row & 127u) * 128u) + (row & 7u);
      a_lo[s] = (uint32_t)addr;
      a_hi[s] = (uint32_t)(addr >> 32);
    }
  };
#pragma unroll
  for (int s = 0; s < kLrMaxSets; ++s) {
    base[s] = r0[s];
    n[s] = 0;
    used[s] = 0;
    cellx[s] = 99;
    wx0[s] = wx1[s] = wx2[s] = wy0[s] = wy1[s] = wy2[s] = 0.f;
    a_lo[s] = a_hi[s] = 0;
    if (s < a.n_sets && r0[s] < r1[s]) refill(s);
  }

  for (;;) {
    // next cell with rows: the smallest head cell over the streams
    int next = 99;
#pragma unroll
    for (int s = 0; s < kLrMaxSets; ++s)
      if (s < a.n_sets && used[s] < n[s]) next = min(next, __shfl_sync(0xffffffffu, cellx[s], used[s]));
    if (next == 99) break;
    if (next - cur >= 3) {          // nothing of the window survives: flush it whole
      slide();
      slide();
      slide();
      cur = next;
    }
    while (cur < next) slide();
#pragma unroll
    for (int s = 0; s < kLrMaxSets; ++s) {
      if (s >= a.n_sets) continue;
      for (;;) {
        if (used[s] >= n[s]) break;
        const unsigned m = __ballot_sync(0xffffffffu, (int)l >= used[s] && (int)l < n[s] && cellx[s] == cur);
        const int cnt = __popc(m);
        if (cnt == 0) break;
        // rows used .. used + cnt - 1 of the batch, four at a time: 8 loads, then 4 x 72 FMAs
        for (int u0 = used[s]; u0 < used[s] + cnt; u0 += 4) {
          uint2 va[4], vb[4];
          float wxv[4][3], wyv[4][3];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const bool on = u0 + u < used[s] + cnt;
            const int src = on ? u0 + u : u0;
            const uint32_t lo = __shfl_sync(0xffffffffu, a_lo[s], src), hi = __shfl_sync(0xffffffffu, a_hi[s], src);
            const uint64_t b64 = ((uint64_t)hi << 32) | (lo & ~127u);
            const uint8_t *p = reinterpret_cast<const uint8_t *>(b64 + (lane_off ^ ((lo & 7u) << 4)));
            va[u] = __ldg(reinterpret_cast<const uint2 *>(p));
            vb[u] = __ldg(reinterpret_cast<const uint2 *>(p + 2 * CN_SLAB_BYTES));
            wxv[u][0] = __shfl_sync(0xffffffffu, wx0[s], src);
            wxv[u][1] = __shfl_sync(0xffffffffu, wx1[s], src);
            wxv[u][2] = __shfl_sync(0xffffffffu, wx2[s], src);
            const float k = on ? 1.f : 0.f;
            wyv[u][0] = k * __shfl_sync(0xffffffffu, wy0[s], src);
            wyv[u][1] = k * __shfl_sync(0xffffffffu, wy1[s], src);
            wyv[u][2] = k * __shfl_sync(0xffffffffu, wy2[s], src);
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const float gv[8] = {bf16_lo(va[u].x), bf16_hi(va[u].x), bf16_lo(va[u].y), bf16_hi(va[u].y),
                                 bf16_lo(vb[u].x), bf16_hi(vb[u].x), bf16_lo(vb[u].y), bf16_hi(vb[u].y)};
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
              for (int ch = 0; ch < 8; ++ch) {
                const float t = wxv[u][c] * gv[ch];
                acc[c][0][ch] = fmaf(wyv[u][0], t, acc[c][0][ch]);
                acc[c][1][ch] = fmaf(wyv[u][1], t, acc[c][1][ch]);
                acc[c][2][ch] = fmaf(wyv[u][2], t, acc[c][2][ch]);
              }
          }
        }
        dirty[0] = dirty[1] = dirty[2] = true;
        used[s] += cnt;
        if (used[s] >= n[s] && base[s] + n[s] < r1[s]) refill(s);   // the cell may continue in the next batch
      }
    }
  }
  slide();
  slide();
  slide();
}

__device__ __forceinline__ void lr_unit_ranges(const LrGradArgs &a, int unit, int &bin, int &ly, int (&r0)[kLrMaxSets],
                                               int (&r1)[kLrMaxSets]) {
  bin = unit / kT;
  ly = unit % kT;
  const int key0 = bin * kCells + ly * kT;
#pragma unroll
  for (int s = 0; s < kLrMaxSets; ++s) {
    r0[s] = r1[s] = 0;
    if (s < a.n_sets) {
      r0[s] = __ldg(a.set[s].key_off + key0);
      r1[s] = key0 + kT < a.n_keys ? __ldg(a.set[s].key_off + key0 + kT) : a.set[s].rows;
    }
  }
}

__global__ void __launch_bounds__(kLrThreads, kLrCtasPerSm)
feature_grad_lr_kernel(CnScene sc, LrGradArgs a, float *__restrict__ d_features) {
  const uint32_t l = lane_id();
  for (;;) {
    int unit = 0;
    if (l == 0) unit = atomicAdd(a.cursor, 1);
    unit = __shfl_sync(0xffffffffu, unit, 0);
    if (unit >= a.n_units) break;
    int bin, ly, r0[kLrMaxSets], r1[kLrMaxSets];
    lr_unit_ranges(a, unit, bin, ly, r0, r1);
    int total = 0, pieces = 0;
#pragma unroll
    for (int s = 0; s < kLrMaxSets; ++s) {
      total += r1[s] - r0[s];
      pieces += (r1[s] - r0[s] + kLrSplit - 1) / kLrSplit;
    }
    if (total == 0) continue;
    if (total > kLrUnitCap) {
      // frame-border cells collect every clamped out-of-frustum sample: split into pieces for the second kernel
      int slot = 0;
      if (l == 0) slot = atomicAdd(a.heavy_n, pieces);
      slot = __shfl_sync(0xffffffffu, slot, 0);
      if (slot + pieces <= a.heavy_cap) {
        int w = slot;
#pragma unroll
        for (int s = 0; s < kLrMaxSets; ++s) {
          const int np = (r1[s] - r0[s] + kLrSplit - 1) / kLrSplit;
          for (int c = (int)l; c < np; c += 32) a.heavy_list[w + c] = make_int2(unit * 4 + s, r0[s] + c * kLrSplit);
          w += np;
        }
        continue;
      }
      // list full (cannot happen with the list sized by heavy_cap_for; kept correct): skip entries, sum here
      for (int c = slot + (int)l; c < min(slot + pieces, a.heavy_cap); c += 32) a.heavy_list[c] = make_int2(-1, 0);
    }
    lr_sweep(sc, a, bin, ly, r0, r1, d_features);
  }
}

__global__ void __launch_bounds__(kLrThreads, kLrCtasPerSm)
feature_grad_lr_heavy_kernel(CnScene sc, LrGradArgs a, float *__restrict__ d_features) {
  const uint32_t l = lane_id();
  const int n_entries = min(a.heavy_n[0], a.heavy_cap);
  for (;;) {
    int e = 0;
    if (l == 0) e = atomicAdd(a.heavy_n + 1, 1);
    e = __shfl_sync(0xffffffffu, e, 0);
    if (e >= n_entries) break;
    const int2 ent = a.heavy_list[e];
    if (ent.x < 0) continue;
    int bin, ly, r0[kLrMaxSets], r1[kLrMaxSets];
    lr_unit_ranges(a, ent.x >> 2, bin, ly, r0, r1);
    const int set = ent.x & 3;
#pragma unroll
    for (int s = 0; s < kLrMaxSets; ++s) {
      if (s == set) {
        r1[s] = min(r1[s], ent.y + kLrSplit);
        r0[s] = ent.y;
      } else {
        r0[s] = r1[s] = 0;
      }
    }
    lr_sweep(sc, a, bin, ly, r0, r1, d_features);
  }
}

}  // namespace cn
