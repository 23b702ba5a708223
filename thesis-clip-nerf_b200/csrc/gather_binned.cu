// Projection of ray samples into the source views, edge-clamped bilinear gather of the 256-channel
// feature map + 3-channel image, positional encoding and assembly of the MLP input rows -- the
// fused replacement of compute_pixel_in_image_mv, get_projection_features_mv,
// world_to_camera_direction_vector_mv, position_encoding and the concat of layers.py:362
// (nerf_utils.py:64-126, 277-285) -- and its backward.  The 259-channel concat copy of the
// reference (nerf_utils.py:278) does not exist here: image and feature map are read in place.
//
// Spatial binning.  A sample-view row reads a 2x2 texel footprint of a [H, W, 256] fp32 map
// (4 KiB) and at 4096 rays x 192 samples x 3 views every texel is touched ~10 times per step, so
// the rows are keyed by (view, 8x8-cell tile, cell) of their clamped footprint, counting-sorted,
// and processed tile by tile:
//   project_bin_kernel   one CTA per 128-row X tile: projection, key, rank (warp-aggregated
//                        atomics), row record; posenc(p), posenc(d), 2*img-1 columns of X staged
//                        in shared memory and written with two 16 KiB bulk stores
//   bin_scan*_kernel     exclusive scan of the key histogram, work items (<= 128 rows of one tile)
//   bin_fill_kernel      records -> sorted order
//   tile_rows_kernel     persistent; a producer warp streams 9x9-texel tiles (81 KiB) of the map
//                        into a 2-deep shared-memory ring with cp.async.bulk + mbarriers, 16
//                        consumer warps serve the rows of the tile from shared memory
//                        (forward: lerp -> bf16 X columns; backward: d/d(pixel) dot products)
//   tile_scatter_kernel  feature-map gradient by PULL: rows are sorted by cell, so each texel of the
//                        tile sums the rows of its (up to) four neighbouring cells in registers
//                        -- no atomics, no shared-memory read-modify-write -- and the finished
//                        tile is added to d_features with cp.reduce.async.bulk.add.f32
//   project_dz_kernel    one CTA per dX tile: posenc'(p), image and pixel-coordinate gradients
//                        chained through K and E^-1 to d/dz (quirk Q4: no stop_gradient)
#include "common.cuh"
#include "project.cuh"

namespace cn {

constexpr int kT = 8;                            // cells per tile edge
constexpr int kTT = kT + 1;                      // texels per tile edge
constexpr int kTexBytes = CN_FEAT_C * 4;         // one texel, fp32
constexpr int kTileBytes = kTT * kTT * kTexBytes;  // 82944
constexpr int kCells = kT * kT;
constexpr int kChunk = 128;                      // max rows per work item

// header ints of the workspace
enum { kHdrItems = 0, kHdrCurFwd = 1, kHdrCurZg = 2, kHdrCurSc = 3, kHdrTicket = 4, kHdrInts = 64 };

struct Ws {
  int64_t rows;
  int tiles_x, tiles_y, n_bins, n_keys, n_blk, max_items;
  int *hdr;        // [kHdrInts]
  int *count;      // [n_keys]   histogram, then (in place) start offset of every key
  int2 *blk_sum;   // [n_blk]    (rows, items) per scan block
  int2 *blk_off;   // [n_blk]    exclusive prefix of blk_sum
  int4 *items;     // [max_items] (bin, start, count, 0)
  int4 *rec_u;     // [rows]     records in row order
  int2 *keyrank;   // [rows]     (key, rank inside key)
  int4 *rec_s;     // [rows]     records sorted by key
  float2 *gpix;    // [rows]     feature part of d loss / d (alpha_x, alpha_y)
  size_t total;
};

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static Ws ws_layout(int b, int nv, int H, int W, int64_t rows, void *base) {
  Ws w;
  w.rows = rows;
  w.tiles_x = (W - 1 + kT - 1) / kT;
  w.tiles_y = (H - 1 + kT - 1) / kT;
  w.n_bins = b * nv * w.tiles_x * w.tiles_y;
  w.n_keys = w.n_bins * kCells;
  w.n_blk = (w.n_bins + 255) / 256;
  w.max_items = (int)(w.n_bins + rows / kChunk + 1);
  uint8_t *p = (uint8_t *)base;
  size_t o = 0;
  auto take = [&](size_t bytes) {
    uint8_t *q = p + o;
    o += align256(bytes);
    return q;
  };
  w.hdr = (int *)take(kHdrInts * 4);
  w.count = (int *)take((size_t)w.n_keys * 4);
  w.blk_sum = (int2 *)take((size_t)w.n_blk * 8);
  w.blk_off = (int2 *)take((size_t)w.n_blk * 8);
  w.items = (int4 *)take((size_t)w.max_items * 16);
  w.rec_u = (int4 *)take((size_t)rows * 16);
  w.keyrank = (int2 *)take((size_t)rows * 8);
  w.rec_s = (int4 *)take((size_t)rows * 16);
  w.gpix = (float2 *)take((size_t)rows * 8);
  w.total = o;
  return w;
}

// ------------------------------------------------------------------------------------------
// projection + key + rank + non-feature columns of X
// ------------------------------------------------------------------------------------------
struct RowId {
  int bv, b, r, i;
  int64_t ray, n;
};
__device__ __forceinline__ RowId decode_row(int64_t row, int nv, int nr, int S) {
  RowId id;
  const int64_t per_bv = (int64_t)nr * S;
  id.bv = (int)(row / per_bv);
  const int64_t rem = row - (int64_t)id.bv * per_bv;
  id.r = (int)(rem / S);
  id.i = (int)(rem - (int64_t)id.r * S);
  id.b = id.bv / nv;
  id.ray = (int64_t)id.b * nr + id.r;
  id.n = id.ray * S + id.i;
  return id;
}

__global__ void __launch_bounds__(128)
project_bin_kernel(CnScene sc, const float *__restrict__ ray_o, const float *__restrict__ ray_d,
                   const float *__restrict__ z, int nr, int S, Ws g, uint8_t *__restrict__ x_tiles) {
  __shared__ __align__(1024) uint8_t sx[2 * CN_SLAB_BYTES];  // slabs 4 and 5 of this X tile
  __shared__ uint32_t sdpe[128][30];                          // posenc(direction) words per ray of the tile
  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int64_t row = (int64_t)blockIdx.x * 128 + tid;
  const RowId id = decode_row(row, sc.nv, nr, S);
  const RowId id0 = decode_row((int64_t)blockIdx.x * 128, sc.nv, nr, S);
  const RayCtx rc = load_ray(ray_o, ray_d, id.ray, z[id.n]);
  const float *E = sc.extr_inv + (int64_t)id.bv * 16;
  const float *K = sc.intrinsics + (int64_t)id.bv * 16;
  const Proj pr = project(E, K, rc.x[0], rc.x[1], rc.x[2], sc.H, sc.W);

  // ---- key, rank inside the key (one atomic per distinct key of the warp), record
  const int lx = pr.x0 & (kT - 1), ly = pr.y0 & (kT - 1);
  const int bin = (id.bv * g.tiles_y + (pr.y0 / kT)) * g.tiles_x + (pr.x0 / kT);
  const int key = bin * kCells + ly * kT + lx;
  {
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const int leader = __ffs(peers) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(g.count + key, __popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    const int rank = base + __popc(peers & ((1u << lane) - 1u));
    g.rec_u[row] = make_int4((int)row, lx | (ly << 8), __float_as_int(pr.ax), __float_as_int(pr.ay));
    g.keyrank[row] = make_int2(key, rank);
  }
  if (!x_tiles) return;

  // ---- posenc(direction): once per (ray, view) of the tile, by the ray's first row in the tile
  const int slot = id.r - id0.r;  // rows of a tile share (b, v); rays are consecutive
  if (id.i == 0 || tid == 0) {
    for (int l = 0; l < 30; ++l) {
      const int c = l / 10, k = l % 10;
      // direction in the source camera frame with homogeneous w = 1 (quirk Q1)
      const float dc = dot4_rn(E + 4 * c, rc.d[0], rc.d[1], rc.d[2], 1.f);
      // libm-accurate sincosf: the values are rounded to bf16 and the parity tests compare ReLU masks
      // downstream of these operands, so the last fp32 bits matter more often than one would think
      float s, co;
      sincosf(__fmul_rn(dc, pe_freq(k)), &s, &co);
      sdpe[slot][l] = pack_bf16x2(s, co);
    }
  }
  // ---- posenc(position): X columns 256..315 = words 0..29 of slab 4
  uint8_t *my4 = sx + tid * 128;
  uint8_t *my5 = sx + CN_SLAB_BYTES + tid * 128;
  const uint32_t sw = (uint32_t)tid & 7u;
  uint32_t w28 = 0, w29 = 0;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int l = q * 4 + j;
      if (l < 30) {
        const int c = l / 10, k = l % 10;
        float s, co;
        sincosf(__fmul_rn(pr.p[c], pe_freq(k)), &s, &co);
        w[j] = pack_bf16x2(s, co);
      } else {
        w[j] = 0;
      }
    }
    if (q < 7) {
      *reinterpret_cast<uint4 *>(my4 + ((q ^ sw) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
    } else {
      w28 = w[0];
      w29 = w[1];
    }
  }
  // ---- 2*img-1, bilinear (X columns 376..378)
  float gi[3];
  {
    const int64_t pix00 = ((int64_t)id.bv * sc.H + pr.y0) * sc.W + pr.x0;
    const float *i00 = sc.images + pix00 * 3;
    const float *i10 = i00 + (int64_t)sc.W * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float tl = __fsub_rn(__fmul_rn(i00[c], 2.f), 1.f), tr = __fsub_rn(__fmul_rn(i00[3 + c], 2.f), 1.f);
      const float bl = __fsub_rn(__fmul_rn(i10[c], 2.f), 1.f), br = __fsub_rn(__fmul_rn(i10[3 + c], 2.f), 1.f);
      gi[c] = lerp2_rn(tl, tr, bl, br, pr.ax, pr.ay);
    }
  }
  __syncthreads();
  // slab 4 chunk 7 = posenc(p) words 28,29 + posenc(d) words 0,1; slab 5 = posenc(d) words 2..29, img, 0
  const uint32_t *dpe = sdpe[slot];
  *reinterpret_cast<uint4 *>(my4 + ((7u ^ sw) << 4)) = make_uint4(w28, w29, dpe[0], dpe[1]);
#pragma unroll
  for (int q = 0; q < 7; ++q)
    *reinterpret_cast<uint4 *>(my5 + ((q ^ sw) << 4)) =
        make_uint4(dpe[2 + 4 * q], dpe[3 + 4 * q], dpe[4 + 4 * q], dpe[5 + 4 * q]);
  *reinterpret_cast<uint4 *>(my5 + ((7u ^ sw) << 4)) = make_uint4(pack_bf16x2(gi[0], gi[1]), pack_bf16x2(gi[2], 0.f), 0u, 0u);
  fence_proxy_async_smem();
  __syncthreads();
  if (tid == 0) {
    uint8_t *dst = x_tiles + (int64_t)blockIdx.x * CN_X_TILE_BYTES + 4 * CN_SLAB_BYTES;
    bulk_s2g(dst, smem_u32(sx), 2 * CN_SLAB_BYTES);
    bulk_commit();
    bulk_wait_read0();
  }
}

// ------------------------------------------------------------------------------------------
// scan of the key histogram and work-item list
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int2 block_excl_scan2(int2 v, int2 *s_warp, int2 &total) {
  // exclusive scan of (x, y) over a 256-thread block
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int2 inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int ax = __shfl_up_sync(0xffffffffu, inc.x, o), ay = __shfl_up_sync(0xffffffffu, inc.y, o);
    if (lane >= o) {
      inc.x += ax;
      inc.y += ay;
    }
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  int2 base = make_int2(0, 0), tot = make_int2(0, 0);
  for (int w = 0; w < 8; ++w) {
    const int2 s = s_warp[w];
    if (w < warp) {
      base.x += s.x;
      base.y += s.y;
    }
    tot.x += s.x;
    tot.y += s.y;
  }
  __syncthreads();
  total = tot;
  return make_int2(base.x + inc.x - v.x, base.y + inc.y - v.y);
}

__device__ __forceinline__ int bin_total(const int *count, int bin) {
  const int4 *c4 = reinterpret_cast<const int4 *>(count + (int64_t)bin * kCells);
  int tot = 0;
#pragma unroll
  for (int q = 0; q < kCells / 4; ++q) {
    const int4 v = c4[q];
    tot += v.x + v.y + v.z + v.w;
  }
  return tot;
}

__global__ void __launch_bounds__(256) bin_scan1_kernel(Ws g) {
  __shared__ int2 s_warp[8];
  __shared__ int s_last;
  const int bin = blockIdx.x * 256 + threadIdx.x;
  const int tot = bin < g.n_bins ? bin_total(g.count, bin) : 0;
  int2 total;
  (void)block_excl_scan2(make_int2(tot, (tot + kChunk - 1) / kChunk), s_warp, total);
  if (threadIdx.x == 0) {
    g.blk_sum[blockIdx.x] = total;
    __threadfence();
    s_last = atomicAdd(g.hdr + kHdrTicket, 1) == (int)gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  // last block: exclusive scan of the block sums
  __threadfence();
  const int per = (g.n_blk + 255) / 256;
  const int i0 = threadIdx.x * per;
  int2 loc = make_int2(0, 0);
  for (int i = i0; i < min(i0 + per, g.n_blk); ++i) {
    const int2 v = __ldcg(g.blk_sum + i);
    loc.x += v.x;
    loc.y += v.y;
  }
  int2 all;
  int2 run = block_excl_scan2(loc, s_warp, all);
  for (int i = i0; i < min(i0 + per, g.n_blk); ++i) {
    const int2 v = __ldcg(g.blk_sum + i);
    g.blk_off[i] = run;
    run.x += v.x;
    run.y += v.y;
  }
  if (threadIdx.x == 0) g.hdr[kHdrItems] = all.y;
}

__global__ void __launch_bounds__(256) bin_scan2_kernel(Ws g) {
  __shared__ int2 s_warp[8];
  const int bin = blockIdx.x * 256 + threadIdx.x;
  const int tot = bin < g.n_bins ? bin_total(g.count, bin) : 0;
  const int nit = (tot + kChunk - 1) / kChunk;
  int2 total;
  const int2 ex = block_excl_scan2(make_int2(tot, nit), s_warp, total);
  if (bin >= g.n_bins) return;
  const int2 off = g.blk_off[blockIdx.x];
  const int row0 = off.x + ex.x, item0 = off.y + ex.y;
  int run = row0;
  int4 *c4 = reinterpret_cast<int4 *>(g.count + (int64_t)bin * kCells);
#pragma unroll 4
  for (int q = 0; q < kCells / 4; ++q) {
    int4 v = c4[q];
    int4 o;
    o.x = run; run += v.x;
    o.y = run; run += v.y;
    o.z = run; run += v.z;
    o.w = run; run += v.w;
    c4[q] = o;
  }
  for (int q = 0; q < nit; ++q) g.items[item0 + q] = make_int4(bin, row0 + q * kChunk, min(kChunk, tot - q * kChunk), 0);
}

__global__ void __launch_bounds__(256) bin_fill_kernel(Ws g) {
  const int64_t row = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (row >= g.rows) return;
  const int2 kr = g.keyrank[row];
  g.rec_s[g.count[kr.x] + kr.y] = g.rec_u[row];
}

// ------------------------------------------------------------------------------------------
// rows of one tile served from shared memory (forward gather / d pixel of the backward)
// ------------------------------------------------------------------------------------------
constexpr int kRowsConsumerWarps = 16;
constexpr int kRowsThreads = 32 * (1 + kRowsConsumerWarps);
constexpr uint32_t kRowsSmem = 2 * kTileBytes + 1024;

struct TileDesc {
  int start, count;
};

// tile origin / extent of a bin
struct TileGeo {
  int bv, x0, y0, nx, ny;
};
__device__ __forceinline__ TileGeo tile_geo(int bin, int tiles_x, int tiles_y, int H, int W) {
  TileGeo t;
  const int tx = bin % tiles_x;
  const int rest = bin / tiles_x;
  const int ty = rest % tiles_y;
  t.bv = rest / tiles_y;
  t.x0 = tx * kT;
  t.y0 = ty * kT;
  t.nx = min(kTT, W - t.x0);
  t.ny = min(kTT, H - t.y0);
  return t;
}

template <bool kBackward>
__global__ void __launch_bounds__(kRowsThreads, 1)
tile_rows_kernel(CnScene sc, Ws g, int *__restrict__ cursor, uint8_t *__restrict__ x_tiles) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[2], bar_empty[2];
  __shared__ TileDesc sdesc[2];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t l = lane_id();

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), kRowsConsumerWarps);
    }
    fence_barrier_init();
  }
  __syncthreads();
  const int n_items = g.hdr[kHdrItems];

  if (warp == 0) {
    // =========================== producer: tiles of the feature map ===========================
    if (l == 0) {
      for (uint32_t k = 0;; ++k) {
        const uint32_t slot = k & 1u;
        mbar_wait(smem_u32(&bar_empty[slot]), ((k >> 1) & 1u) ^ 1u);
        const int it = atomicAdd(cursor, 1);
        if (it >= n_items) {
          sdesc[slot].count = 0;
          mbar_arrive(smem_u32(&bar_full[slot]));
          break;
        }
        const int4 item = g.items[it];
        const TileGeo t = tile_geo(item.x, g.tiles_x, g.tiles_y, sc.H, sc.W);
        sdesc[slot].start = item.y;
        sdesc[slot].count = item.z;
        const uint32_t row_bytes = (uint32_t)t.nx * kTexBytes;
        mbar_expect_tx(smem_u32(&bar_full[slot]), row_bytes * (uint32_t)t.ny);
        const float *src = sc.features + (((int64_t)t.bv * sc.H + t.y0) * sc.W + t.x0) * CN_FEAT_C;
        const uint32_t dst = smem_u32(smem) + slot * kTileBytes;
        for (int yy = 0; yy < t.ny; ++yy)
          bulk_g2s(dst + yy * (kTT * kTexBytes), src + (int64_t)yy * sc.W * CN_FEAT_C, row_bytes,
                   smem_u32(&bar_full[slot]));
      }
    }
  } else {
    // =========================== consumers: one warp per row ===========================
    const uint32_t cw = warp - 1;
    for (uint32_t k = 0;; ++k) {
      const uint32_t slot = k & 1u;
      mbar_wait(smem_u32(&bar_full[slot]), (k >> 1) & 1u);
      const int start = sdesc[slot].start, count = sdesc[slot].count;
      if (count == 0) break;
      const float *tile = reinterpret_cast<const float *>(smem + slot * kTileBytes);
      int4 rec = make_int4(0, 0, 0, 0);
      uint2 ga = make_uint2(0, 0), gb = make_uint2(0, 0);
      auto fetch = [&](int j) {
        rec = __ldg(g.rec_s + start + j);
        if (kBackward) {
          const uint8_t *t = x_tiles + (int64_t)(rec.x >> 7) * CN_X_TILE_BYTES;
          const uint32_t rr = (uint32_t)rec.x & 127u;
          ga = __ldg(reinterpret_cast<const uint2 *>(t + tile_chunk_off(rr, l >> 1) + (l & 1) * 8));
          gb = __ldg(reinterpret_cast<const uint2 *>(t + tile_chunk_off(rr, 16 + (l >> 1)) + (l & 1) * 8));
        }
      };
      if ((int)cw < count) fetch(cw);
      for (int j = cw; j < count; j += kRowsConsumerWarps) {
        const int4 cur = rec;
        const uint2 ca = ga, cb = gb;
        if (j + kRowsConsumerWarps < count) fetch(j + kRowsConsumerWarps);
        const int row = cur.x;
        const int lx = cur.y & 0xff, ly = (cur.y >> 8) & 0xff;
        const float ax = __int_as_float(cur.z), ay = __int_as_float(cur.w);
        const float *t00 = tile + (ly * kTT + lx) * CN_FEAT_C;
        const float *t10 = t00 + kTT * CN_FEAT_C;
        // channels 4l..4l+3 and 128+4l..128+4l+3 of the four corners (conflict-free 512-byte reads)
        const float4 tl0 = *reinterpret_cast<const float4 *>(t00 + 4 * l);
        const float4 tl1 = *reinterpret_cast<const float4 *>(t00 + 128 + 4 * l);
        const float4 tr0 = *reinterpret_cast<const float4 *>(t00 + CN_FEAT_C + 4 * l);
        const float4 tr1 = *reinterpret_cast<const float4 *>(t00 + CN_FEAT_C + 128 + 4 * l);
        const float4 bl0 = *reinterpret_cast<const float4 *>(t10 + 4 * l);
        const float4 bl1 = *reinterpret_cast<const float4 *>(t10 + 128 + 4 * l);
        const float4 br0 = *reinterpret_cast<const float4 *>(t10 + CN_FEAT_C + 4 * l);
        const float4 br1 = *reinterpret_cast<const float4 *>(t10 + CN_FEAT_C + 128 + 4 * l);
        if (!kBackward) {
          uint8_t *xt = x_tiles + (int64_t)(row >> 7) * CN_X_TILE_BYTES;
          const uint32_t rr = (uint32_t)row & 127u;
          const float4 g0 = lerp2_f4(tl0, tr0, bl0, br0, ax, ay);
          const float4 g1 = lerp2_f4(tl1, tr1, bl1, br1, ax, ay);
          // channels 4l..4l+3 -> half of chunk l/2; channels 128+4l.. -> half of chunk 16 + l/2
          *reinterpret_cast<uint2 *>(xt + tile_chunk_off(rr, l >> 1) + (l & 1) * 8) =
              make_uint2(pack_bf16x2(g0.x, g0.y), pack_bf16x2(g0.z, g0.w));
          *reinterpret_cast<uint2 *>(xt + tile_chunk_off(rr, 16 + (l >> 1)) + (l & 1) * 8) =
              make_uint2(pack_bf16x2(g1.x, g1.y), pack_bf16x2(g1.z, g1.w));
        } else {
          // d out / d ax = (1-ay)(TR-TL) + ay(BR-BL);  d out / d ay = (1-ax)(BL-TL) + ax(BR-TR)
          const float g0[4] = {bf16_lo(ca.x), bf16_hi(ca.x), bf16_lo(ca.y), bf16_hi(ca.y)};
          const float g1[4] = {bf16_lo(cb.x), bf16_hi(cb.x), bf16_lo(cb.y), bf16_hi(cb.y)};
          const float tlv[8] = {tl0.x, tl0.y, tl0.z, tl0.w, tl1.x, tl1.y, tl1.z, tl1.w};
          const float trv[8] = {tr0.x, tr0.y, tr0.z, tr0.w, tr1.x, tr1.y, tr1.z, tr1.w};
          const float blv[8] = {bl0.x, bl0.y, bl0.z, bl0.w, bl1.x, bl1.y, bl1.z, bl1.w};
          const float brv[8] = {br0.x, br0.y, br0.z, br0.w, br1.x, br1.y, br1.z, br1.w};
          float gax = 0.f, gay = 0.f;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const float gc = c < 4 ? g0[c] : g1[c - 4];
            gax += gc * ((1.f - ay) * (trv[c] - tlv[c]) + ay * (brv[c] - blv[c]));
            gay += gc * ((1.f - ax) * (blv[c] - tlv[c]) + ax * (brv[c] - trv[c]));
          }
          gax = warp_sum_f(gax);
          gay = warp_sum_f(gay);
          if (l == 0) g.gpix[row] = make_float2(gax, gay);
        }
      }
      __syncwarp();
      if (l == 0) mbar_arrive(smem_u32(&bar_empty[slot]));
    }
  }
}

// ------------------------------------------------------------------------------------------
// feature-map gradient: pull-accumulate per texel, bulk reduce-add per tile row
// ------------------------------------------------------------------------------------------
constexpr int kScatterWarps = 8;
constexpr int kScatterThreads = 32 * kScatterWarps;
constexpr uint32_t kScatterSmem = kTileBytes + 1024;

struct SRec {
  uint32_t base128;  // (row >> 7) * 768 + (row & 127): start of the row inside the dX image, in 128-byte units
  uint32_t cell_sw;  // cell | (row & 7) << 8
  float ax, ay;
};

__device__ __forceinline__ SRec make_srec(int4 r) {
  SRec s;
  const uint32_t row = (uint32_t)r.x;
  s.base128 = (row >> 7) * (CN_X_TILE_BYTES / 128) + (row & 127u);
  const uint32_t lx = (uint32_t)r.y & 0xffu, ly = ((uint32_t)r.y >> 8) & 0xffu;
  s.cell_sw = (ly * kT + lx) | ((row & 7u) << 8);
  s.ax = __int_as_float(r.z);
  s.ay = __int_as_float(r.w);
  return s;
}

__global__ void __launch_bounds__(kScatterThreads, 2)
tile_scatter_kernel(CnScene sc, Ws g, int *__restrict__ cursor, const uint8_t *__restrict__ dx_tiles,
                    float *__restrict__ d_features) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(16) SRec srec[2][kChunk];
  __shared__ int cstart[kCells + 1];
  __shared__ int s_claim[2];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  float *out = reinterpret_cast<float *>(smem);
  const int tid = threadIdx.x;
  const uint32_t warp = tid >> 5, l = lane_id();
  const int n_items = g.hdr[kHdrItems];

  if (tid == 0) {
    s_claim[0] = atomicAdd(cursor, 1);
    s_claim[1] = atomicAdd(cursor, 1);
  }
  __syncthreads();
  int cur = s_claim[0], nxt = s_claim[1];
  __syncthreads();
  int4 item = make_int4(0, 0, 0, 0);
  if (cur < n_items) {
    item = g.items[cur];
    if (tid < item.z) srec[0][tid] = make_srec(__ldg(g.rec_s + item.y + tid));
  }
  __syncthreads();
  bool flush_pending = false;
  // lane l owns bf16 chunk l of a dX row = channels 8l..8l+7: slab l/8, 16-byte chunk (l%8) ^ (row%8)
  const uint32_t lane_slab_off = (l >> 3) * CN_SLAB_BYTES;

  for (uint32_t k = 0; cur < n_items; ++k) {
    const uint32_t buf = k & 1u;
    // ---- prefetch the next item's records, claim the one after
    int4 item_n = make_int4(0, 0, 0, 0), rec_n = make_int4(0, 0, 0, 0);
    int claim = 0;
    if (nxt < n_items) {
      item_n = g.items[nxt];
      if (tid < item_n.z) rec_n = __ldg(g.rec_s + item_n.y + tid);
    }
    if (tid == 0) claim = atomicAdd(cursor, 1);
    // ---- first row of every cell (rows of an item are sorted by cell)
    const int count = item.z;
    if (tid < count) {
      const int c = (int)(srec[buf][tid].cell_sw & 0xffu);
      const int cp = tid > 0 ? (int)(srec[buf][tid - 1].cell_sw & 0xffu) : -1;
      for (int q = cp + 1; q <= c; ++q) cstart[q] = tid;
      if (tid == count - 1)
        for (int q = c + 1; q <= kCells; ++q) cstart[q] = count;
    }
    if (flush_pending && tid == 0) bulk_wait_read0();  // the previous tile has left shared memory
    __syncthreads();

    const TileGeo t = tile_geo(item.x, g.tiles_x, g.tiles_y, sc.H, sc.W);
    for (int tex = warp; tex < kTT * kTT; tex += kScatterWarps) {
      const int y = tex / kTT, x = tex - y * kTT;
      if (y >= t.ny || x >= t.nx) continue;
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int corner = 0; corner < 4; ++corner) {
        const int dy = corner >> 1, dx = corner & 1;
        const int cy = y - dy, cx = x - dx;
        if (cy < 0 || cy >= kT || cx < 0 || cx >= kT) continue;
        const int c = cy * kT + cx;
        const int j1 = cstart[c + 1];
        for (int j = cstart[c]; j < j1; j += 4) {
          // four independent 16-byte loads in flight per lane before the FMAs
          uint4 v[4];
          float w[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (j + u < j1) {
              const SRec r = srec[buf][j + u];
              const float wx = dx ? r.ax : 1.f - r.ax;
              const float wy = dy ? r.ay : 1.f - r.ay;
              w[u] = wx * wy;
              const uint8_t *p =
                  dx_tiles + (int64_t)r.base128 * 128 + lane_slab_off + (((l & 7u) ^ (r.cell_sw >> 8)) << 4);
              v[u] = __ldg(reinterpret_cast<const uint4 *>(p));
            } else {
              w[u] = 0.f;
              v[u] = make_uint4(0u, 0u, 0u, 0u);
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            acc[0] = fmaf(w[u], bf16_lo(v[u].x), acc[0]); acc[1] = fmaf(w[u], bf16_hi(v[u].x), acc[1]);
            acc[2] = fmaf(w[u], bf16_lo(v[u].y), acc[2]); acc[3] = fmaf(w[u], bf16_hi(v[u].y), acc[3]);
            acc[4] = fmaf(w[u], bf16_lo(v[u].z), acc[4]); acc[5] = fmaf(w[u], bf16_hi(v[u].z), acc[5]);
            acc[6] = fmaf(w[u], bf16_lo(v[u].w), acc[6]); acc[7] = fmaf(w[u], bf16_hi(v[u].w), acc[7]);
          }
        }
      }
      float4 *o = reinterpret_cast<float4 *>(out + tex * CN_FEAT_C + 8 * l);
      o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
      o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
      float *dst = d_features + (((int64_t)t.bv * sc.H + t.y0) * sc.W + t.x0) * CN_FEAT_C;
      for (int yy = 0; yy < t.ny; ++yy)
        bulk_reduce_add_f32(dst + (int64_t)yy * sc.W * CN_FEAT_C, smem_u32(out) + yy * (kTT * kTexBytes),
                            (uint32_t)t.nx * kTexBytes);
      bulk_commit();
      s_claim[0] = claim;
    }
    flush_pending = true;
    if (nxt < n_items && tid < item_n.z) srec[buf ^ 1u][tid] = make_srec(rec_n);
    __syncthreads();
    cur = nxt;
    nxt = s_claim[0];
    item = item_n;
  }
  if (flush_pending && tid == 0) bulk_wait0();
}

// ------------------------------------------------------------------------------------------
// d/dz: posenc'(p), image and pixel-coordinate gradients chained through K and E^-1
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
project_dz_kernel(CnScene sc, const float *__restrict__ ray_o, const float *__restrict__ ray_d,
                  const float *__restrict__ z, int nr, int S, const uint8_t *__restrict__ dx_tiles,
                  const float2 *__restrict__ gpix, float *__restrict__ dz) {
  __shared__ __align__(1024) uint8_t sx[2 * CN_SLAB_BYTES];  // slabs 4 and 5 of this dX tile
  __shared__ __align__(8) uint64_t bar;
  const int tid = threadIdx.x;
  if (tid == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
    mbar_expect_tx(smem_u32(&bar), 2 * CN_SLAB_BYTES);
    bulk_g2s(smem_u32(sx), dx_tiles + (int64_t)blockIdx.x * CN_X_TILE_BYTES + 4 * CN_SLAB_BYTES, 2 * CN_SLAB_BYTES,
             smem_u32(&bar));
  }
  const int64_t row = (int64_t)blockIdx.x * 128 + tid;
  const RowId id = decode_row(row, sc.nv, nr, S);
  const RayCtx rc = load_ray(ray_o, ray_d, id.ray, z[id.n]);
  const float *E = sc.extr_inv + (int64_t)id.bv * 16;
  const float *K = sc.intrinsics + (int64_t)id.bv * 16;
  const Proj pr = project(E, K, rc.x[0], rc.x[1], rc.x[2], sc.H, sc.W);
  const float2 gf = gpix[row];
  float gax = gf.x, gay = gf.y;
  __syncthreads();
  mbar_wait(smem_u32(&bar), 0);

  const uint8_t *my4 = sx + tid * 128;
  const uint8_t *my5 = sx + CN_SLAB_BYTES + tid * 128;
  const uint32_t sw = (uint32_t)tid & 7u;
  // image columns 376..378 = slab 5, words 28, 29
  {
    const uint4 gi = *reinterpret_cast<const uint4 *>(my5 + ((7u ^ sw) << 4));
    const float gimg[3] = {bf16_lo(gi.x), bf16_hi(gi.x), bf16_lo(gi.y)};
    const int64_t pix00 = ((int64_t)id.bv * sc.H + pr.y0) * sc.W + pr.x0;
    const float *i00 = sc.images + pix00 * 3;
    const float *i10 = i00 + (int64_t)sc.W * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float tl = 2.f * i00[c], tr = 2.f * i00[3 + c], bl = 2.f * i10[c], br = 2.f * i10[3 + c];
      gax += gimg[c] * ((1.f - pr.ay) * (tr - tl) + pr.ay * (br - bl));
      gay += gimg[c] * ((1.f - pr.ax) * (bl - tl) + pr.ax * (br - tr));
    }
  }
  // posenc(position): d/dp_c sin(p_c f_k) = f_k cos, d/dp_c cos(p_c f_k) = -f_k sin
  float dp[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const uint4 v = *reinterpret_cast<const uint4 *>(my4 + ((q ^ sw) << 4));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int lw = q * 4 + j;
      if (lw < 30) {
        const int c = lw / 10, k = lw % 10;
        const float fk = pe_freq(k);
        float s, co;
        sincos_posenc(__fmul_rn(pr.p[c], fk), s, co);
        dp[c] += fk * (bf16_lo(w[j]) * co - bf16_hi(w[j]) * s);
      }
    }
  }
  // alpha clamp, +-1e6 clip and the max(q.z, 1e-8) branch gate the coordinate gradient
  const float gpx = (pr.tx >= 0.f && pr.tx <= 1.f && pr.rx >= -1e6f && pr.rx <= 1e6f) ? gax : 0.f;
  const float gpy = (pr.ty >= 0.f && pr.ty <= 1.f && pr.ry >= -1e6f && pr.ry <= 1e6f) ? gay : 0.f;
  const float dq0 = gpx / pr.den, dq1 = gpy / pr.den;
  const float dq2 = (pr.q[2] >= 1e-8f) ? -(gpx * pr.rx + gpy * pr.ry) / pr.den : 0.f;
#pragma unroll
  for (int j = 0; j < 3; ++j) dp[j] += K[0 + j] * dq0 + K[4 + j] * dq1 + K[8 + j] * dq2;
  // p = E [o + z d; 1]  ->  dp_i/dz = sum_j E_ij d_j
  float acc = 0.f;
#pragma unroll
  for (int ii = 0; ii < 3; ++ii)
    acc += dp[ii] * (E[4 * ii + 0] * rc.d[0] + E[4 * ii + 1] * rc.d[1] + E[4 * ii + 2] * rc.d[2]);
  atomicAdd(dz + id.n, acc);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static int check_rows(const CnScene *scene, int nr, int S, const char *who, int64_t *rows) {
  CN_REQUIRE(nr > 0 && S > 0 && ((int64_t)nr * S) % CN_TILE_ROWS == 0,
             "%s: nr*S must be a positive multiple of %d (nr=%d S=%d)", who, CN_TILE_ROWS, nr, S);
  *rows = (int64_t)scene->b * scene->nv * nr * S;
  CN_REQUIRE(*rows < ((int64_t)1 << 31), "%s: %lld sample-view rows exceed 2^31", who, (long long)*rows);
  CN_REQUIRE((int64_t)scene->b * scene->nv * ((scene->W + kT - 2) / kT) * ((scene->H + kT - 2) / kT) * kCells <
                 ((int64_t)1 << 31),
             "%s: too many bins", who);
  return 0;
}

static int get_ws(const CnScene *scene, int nr, int S, void *ws, size_t ws_bytes, const char *who, Ws *out) {
  int64_t rows;
  if (int e = check_rows(scene, nr, S, who, &rows)) return e;
  CN_REQUIRE(ws && ((uintptr_t)ws & 255) == 0, "%s: workspace must be a 256-byte aligned device pointer", who);
  *out = ws_layout(scene->b, scene->nv, scene->H, scene->W, rows, ws);
  CN_REQUIRE(ws_bytes >= out->total, "%s: workspace too small (%zu < %zu bytes)", who, ws_bytes, out->total);
  return 0;
}

static int run_bin(const CnScene *scene, const float *ray_o, const float *ray_d, const float *z, int nr, int S,
                   const Ws &g, uint8_t *x_tiles, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(g.hdr, 0, (size_t)((uint8_t *)g.blk_sum - (uint8_t *)g.hdr), st);  // hdr + count
  if (e != cudaSuccess) {
    set_error("cn_project_bin: cudaMemsetAsync: %s", cudaGetErrorString(e));
    return (int)e;
  }
  project_bin_kernel<<<(unsigned)(g.rows / 128), 128, 0, st>>>(*scene, ray_o, ray_d, z, nr, S, g, x_tiles);
  bin_scan1_kernel<<<g.n_blk, 256, 0, st>>>(g);
  bin_scan2_kernel<<<g.n_blk, 256, 0, st>>>(g);
  bin_fill_kernel<<<(unsigned)((g.rows + 255) / 256), 256, 0, st>>>(g);
  return check_launch("cn_project_bin");
}

template <bool kBackward>
static int run_tile_rows(const CnScene *scene, const Ws &g, int cursor_index, uint8_t *x_tiles, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(tile_rows_kernel<kBackward>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kRowsSmem);
  if (e != cudaSuccess) {
    set_error("tile_rows: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  tile_rows_kernel<kBackward><<<kNumSMs, kRowsThreads, kRowsSmem, st>>>(*scene, g, g.hdr + cursor_index, x_tiles);
  return check_launch("tile_rows");
}

}  // namespace cn

extern "C" size_t cn_gather_ws_bytes(int32_t b, int32_t nv, int32_t H, int32_t W, int32_t nr, int32_t S) {
  if (b <= 0 || nv <= 0 || H < 2 || W < 2 || nr <= 0 || S <= 0) return 0;
  return cn::ws_layout(b, nv, H, W, (int64_t)b * nv * nr * S, nullptr).total;
}

extern "C" int cn_project_bin(const CnScene *scene, const float *ray_o, const float *ray_d, const float *z,
                              int32_t nr, int32_t S, void *ws, size_t ws_bytes, void *stream) {
  using namespace cn;
  if (int e = check_scene(scene, "cn_project_bin")) return e;
  CN_REQUIRE(ray_o && ray_d && z, "cn_project_bin: null pointer");
  Ws g;
  if (int e = get_ws(scene, nr, S, ws, ws_bytes, "cn_project_bin", &g)) return e;
  return run_bin(scene, ray_o, ray_d, z, nr, S, g, nullptr, (cudaStream_t)stream);
}

extern "C" int cn_project_gather_fwd(const CnScene *scene, const float *ray_o, const float *ray_d, const float *z,
                                     int32_t nr, int32_t S, void *x_tiles, void *ws, size_t ws_bytes, void *stream) {
  using namespace cn;
  if (int e = check_scene(scene, "cn_project_gather_fwd")) return e;
  CN_REQUIRE(ray_o && ray_d && z && x_tiles, "cn_project_gather_fwd: null pointer");
  CN_REQUIRE(((uintptr_t)x_tiles & 127) == 0, "cn_project_gather_fwd: x_tiles must be 128-byte aligned");
  Ws g;
  if (int e = get_ws(scene, nr, S, ws, ws_bytes, "cn_project_gather_fwd", &g)) return e;
  if (int e = run_bin(scene, ray_o, ray_d, z, nr, S, g, (uint8_t *)x_tiles, (cudaStream_t)stream)) return e;
  return run_tile_rows<false>(scene, g, kHdrCurFwd, (uint8_t *)x_tiles, (cudaStream_t)stream);
}

extern "C" int cn_project_gather_bwd(const CnScene *scene, const float *ray_o, const float *ray_d, const float *z,
                                     int32_t nr, int32_t S, const void *dx_tiles, void *ws, size_t ws_bytes,
                                     float *d_features, float *dz, void *stream) {
  using namespace cn;
  if (int e = check_scene(scene, "cn_project_gather_bwd")) return e;
  CN_REQUIRE(ray_o && ray_d && z && dx_tiles, "cn_project_gather_bwd: null pointer");
  CN_REQUIRE(((uintptr_t)dx_tiles & 127) == 0, "cn_project_gather_bwd: dx_tiles must be 128-byte aligned");
  CN_REQUIRE(!d_features || ((uintptr_t)d_features & 15) == 0, "cn_project_gather_bwd: d_features alignment");
  Ws g;
  if (int e = get_ws(scene, nr, S, ws, ws_bytes, "cn_project_gather_bwd", &g)) return e;
  if (!d_features && !dz) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t ce = cudaMemsetAsync(g.hdr + kHdrCurZg, 0, 2 * sizeof(int), st);  // both backward cursors
  if (ce != cudaSuccess) {
    set_error("cn_project_gather_bwd: cudaMemsetAsync: %s", cudaGetErrorString(ce));
    return (int)ce;
  }
  if (dz) {
    if (int e = run_tile_rows<true>(scene, g, kHdrCurZg, (uint8_t *)dx_tiles, st)) return e;
    project_dz_kernel<<<(unsigned)(g.rows / 128), 128, 0, st>>>(*scene, ray_o, ray_d, z, nr, S,
                                                               (const uint8_t *)dx_tiles, g.gpix, dz);
    if (int e = check_launch("cn_project_gather_bwd (dz)")) return e;
  }
  if (d_features) {
    ce = cudaFuncSetAttribute(tile_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kScatterSmem);
    if (ce != cudaSuccess) {
      set_error("cn_project_gather_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(ce));
      return (int)ce;
    }
    tile_scatter_kernel<<<2 * kNumSMs, kScatterThreads, kScatterSmem, st>>>(*scene, g, g.hdr + kHdrCurSc,
                                                                            (const uint8_t *)dx_tiles, d_features);
    if (int e = check_launch("cn_project_gather_bwd (scatter)")) return e;
  }
  return 0;
}
