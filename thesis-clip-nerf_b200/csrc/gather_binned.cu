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
#include <cstdlib>

#include "common.cuh"
#include "project.cuh"

namespace cn {

constexpr int kT = 8;                            // cells per tile edge
constexpr int kTT = kT + 1;                      // texels per tile edge
constexpr int kTexBytes = CN_FEAT_C * 4;         // one texel, fp32
constexpr int kTileBytes = kTT * kTT * kTexBytes;  // 82944
constexpr int kCells = kT * kT;
#ifndef CN_CHUNK
#define CN_CHUNK 128
#endif
constexpr int kChunk = CN_CHUNK;                 // max rows per work item

// header ints of the workspace
enum { kHdrItems = 0, kHdrCurFwd = 1, kHdrCurZg = 2, kHdrCurSc = 3, kHdrHeavyN = 4 /* 2 ints */, kHdrTicket = 6, kHdrInts = 64 };
// Deferred (texel, chunk) entries of the feature-gradient sweep.  A deferred texel has n > kHeavyRows row visits and
// ceil(n / kHeavyChunk) <= n (1/256 + 1/193) entries, and every row is visited by 4 texels, so all sets of one sweep
// (<= 4, each at most as large as the largest) need at most 4 * 4 * rows_max * 0.0091 < rows_max / 4 entries: the list
// is sized from the row count of its workspace and cn_feature_grad takes the list of the LARGEST set -- it cannot fill
// up.  (The sweep still handles a full list -- see feature_grad_kernel -- which tests reach through CN_DEBUG_HEAVY_CAP.)
static int heavy_cap_for(int64_t rows) { return (int)(rows / 4 + 1024); }

constexpr int kChunkLr = 256;                    // feat_up == 2: rows per work item (its tile is 100 KiB, see gather_lowres.cuh)

struct Ws {
  int64_t rows;
  int tiles_x, tiles_y, n_bins, n_keys, n_blk, max_items, heavy_cap;
  int chunk;       // max rows per work item
  int *hdr;        // [kHdrInts]
  int *count;      // [n_keys]   histogram, then (in place) start offset of every key
  int2 *blk_sum;   // [n_blk]    (rows, items) per scan block
  int2 *blk_off;   // [n_blk]    exclusive prefix of blk_sum
  int4 *items;     // [max_items] (bin, start, count, 0)
  int4 *rec_u;     // [rows]     records in row order
  int2 *keyrank;   // [rows]     (key, rank inside key)
  int4 *rec_s;     // [rows]     records sorted by key
  float2 *gpix;    // [rows]     feature part of d loss / d (alpha_x, alpha_y)
  int2 *heavy;     // [heavy_cap] (texel, first row) chunks deferred by feature_grad_kernel
  size_t total;
};

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// Cells of the key grid: feat_up == 1, the (W-1) x (H-1) clamped 2x2 footprints of the map itself; feat_up == 2, the
// (W/2) x (H/2) base texels k = f0 >> 1 of the composed stencil on the low-resolution map (f0 <= W - 2, so k <= W/2 - 1).
// The feat_up == 2 layout is never larger than the feat_up == 1 one: cn_gather_ws_bytes bounds both.
static Ws ws_layout(int b, int nv, int H, int W, int up, int64_t rows, void *base) {
  Ws w;
  w.rows = rows;
  const int cells_x = up == 2 ? W / 2 : W - 1, cells_y = up == 2 ? H / 2 : H - 1;
  w.chunk = up == 2 ? kChunkLr : kChunk;
  w.tiles_x = (cells_x + kT - 1) / kT;
  w.tiles_y = (cells_y + kT - 1) / kT;
  w.n_bins = b * nv * w.tiles_x * w.tiles_y;
  w.n_keys = w.n_bins * kCells;
  w.n_blk = (w.n_bins + 255) / 256;
  w.max_items = (int)(w.n_bins + rows / w.chunk + 1);
  w.heavy_cap = heavy_cap_for(rows);
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
  w.heavy = (int2 *)take((size_t)w.heavy_cap * 8);
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
  extern __shared__ uint32_t sdpe_dyn[];                      // posenc(direction) words, [rays of the tile][30]
  uint32_t(*sdpe)[30] = reinterpret_cast<uint32_t(*)[30]>(sdpe_dyn);
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
  // feat_up == 2: the cell is the base texel (x0 >> 1, y0 >> 1) of the low-resolution stencil; the parities of x0, y0
  // select its weights and ride in bits 16, 17 of the record's cell word
  const int up_shift = sc.feat_up == 2 ? 1 : 0;
  const int cx = pr.x0 >> up_shift, cy = pr.y0 >> up_shift;
  const int par = up_shift ? ((pr.x0 & 1) | ((pr.y0 & 1) << 1)) : 0;
  const int lx = cx & (kT - 1), ly = cy & (kT - 1);
  const int bin = (id.bv * g.tiles_y + (cy / kT)) * g.tiles_x + (cx / kT);
  const int key = bin * kCells + ly * kT + lx;
  {
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const int leader = __ffs(peers) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(g.count + key, __popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    const int rank = base + __popc(peers & ((1u << lane) - 1u));
    g.rec_u[row] = make_int4((int)row, lx | (ly << 8) | (par << 16), __float_as_int(pr.ax), __float_as_int(pr.ay));
    g.keyrank[row] = make_int2(key, rank);
  }
  if (!x_tiles) return;

  // ---- posenc(direction): once per (ray, view) of the tile, its 30 (coordinate, frequency) pairs spread
  // over the rows the ray has in this tile
  const int seg0 = max(0, tid - id.i);                 // first thread of this ray's segment
  const int seg_len = min(128 - seg0, S - (id.i - (tid - seg0)));
  const int slot = id.r - id0.r;                       // rows of a tile share (b, v); its rays are consecutive
  for (int l = tid - seg0; l < 30; l += seg_len) {
    const int c = l / 10, k = l % 10;
    // direction in the source camera frame with homogeneous w = 1 (quirk Q1)
    const float dc = dot4_rn(E + 4 * c, rc.d[0], rc.d[1], rc.d[2], 1.f);
    // libm-accurate sincosf: the values are rounded to bf16 and the parity tests compare ReLU masks
    // downstream of these operands, so the last fp32 bits matter more often than one would think
    float s, co;
    sincosf(__fmul_rn(dc, pe_freq(k)), &s, &co);
    sdpe[slot][l] = pack_bf16x2(s, co);
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
  (void)block_excl_scan2(make_int2(tot, (tot + g.chunk - 1) / g.chunk), s_warp, total);
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
  const int nit = (tot + g.chunk - 1) / g.chunk;
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
  for (int q = 0; q < nit; ++q) g.items[item0 + q] = make_int4(bin, row0 + q * g.chunk, min(g.chunk, tot - q * g.chunk), 0);
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
// consumer warps per CTA (one CTA per SM: the two tile slots fill its shared memory).  Measured (B200, cfg2): 24 / 19
// and 31 / 23 consumer warps (forward / backward) are 1 - 4 % SLOWER than 16 / 16 -- the kernels wait for tiles, not
// for issue slots or warps
#ifndef CN_ROWS_WARPS_FWD
#define CN_ROWS_WARPS_FWD 16
#endif
#ifndef CN_ROWS_WARPS_BWD
#define CN_ROWS_WARPS_BWD 16
#endif
// L2 prefetch of the next work item's tile while both slots are busy: measured -3 % on the backward sweep but +12 %
// on the forward one (it competes with the X stores for L2); off
#ifndef CN_DIAG
#define CN_DIAG 0     // 1 / 2 / 3: timing diagnostics of the tile sweeps (wrong results): no tile loads / no row work / no dX loads
#endif
#ifndef CN_TILE_L2_PREFETCH
#define CN_TILE_L2_PREFETCH 0
#endif
template <bool kBackward>
constexpr int rows_consumer_warps() { return kBackward ? CN_ROWS_WARPS_BWD : CN_ROWS_WARPS_FWD; }
template <bool kBackward>
constexpr int rows_threads() { return 32 * (1 + rows_consumer_warps<kBackward>()); }
constexpr uint32_t kRowsRecBytes = kChunk * 16;                       // records of one work item
constexpr uint32_t kRowsSlotBytes = kTileBytes + kRowsRecBytes;       // tile + records
constexpr uint32_t kRowsSmem = 2 * kRowsSlotBytes + 1024;

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

struct Corners {
  float4 tl0, tl1, tr0, tr1, bl0, bl1, br0, br1;
};
// channels 4l..4l+3 and 128+4l..128+4l+3 of the four corners (conflict-free 512-byte reads)
__device__ __forceinline__ Corners load_corners(const float *tile, int loc, uint32_t l) {
  const int lx = loc & 0xff, ly = (loc >> 8) & 0xff;
  const float *t00 = tile + (ly * kTT + lx) * CN_FEAT_C;
  const float *t10 = t00 + kTT * CN_FEAT_C;
  Corners c;
  c.tl0 = *reinterpret_cast<const float4 *>(t00 + 4 * l);
  c.tl1 = *reinterpret_cast<const float4 *>(t00 + 128 + 4 * l);
  c.tr0 = *reinterpret_cast<const float4 *>(t00 + CN_FEAT_C + 4 * l);
  c.tr1 = *reinterpret_cast<const float4 *>(t00 + CN_FEAT_C + 128 + 4 * l);
  c.bl0 = *reinterpret_cast<const float4 *>(t10 + 4 * l);
  c.bl1 = *reinterpret_cast<const float4 *>(t10 + 128 + 4 * l);
  c.br0 = *reinterpret_cast<const float4 *>(t10 + CN_FEAT_C + 4 * l);
  c.br1 = *reinterpret_cast<const float4 *>(t10 + CN_FEAT_C + 128 + 4 * l);
  return c;
}

template <bool kBackward>
__global__ void __launch_bounds__(rows_threads<kBackward>(), 1)
tile_rows_kernel(CnScene sc, Ws g, int *__restrict__ cursor, uint8_t *__restrict__ x_tiles) {
  constexpr int kRowsConsumerWarps = rows_consumer_warps<kBackward>();
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
    // ============ producer: 9x9-texel tile of the feature map + the item's row records ============
    if (l == 0) {
      // the descriptor of the NEXT work item (cursor atomic + dependent load, ~1.5 us of latency) is fetched right
      // after the copies of the current one are issued, not after its slot is released
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
        const TileGeo t = tile_geo(item.x, g.tiles_x, g.tiles_y, sc.H, sc.W);
        sdesc[slot].start = item.y;
        sdesc[slot].count = item.z;
        const uint32_t row_bytes = (uint32_t)t.nx * kTexBytes;
        const uint32_t bar = smem_u32(&bar_full[slot]);
#if CN_DIAG == 1   // timing diagnostic only (tools/diag_tile_rows.sh): no tile loads, the rows read stale shared memory
        mbar_expect_tx(bar, (uint32_t)item.z * 16u);
#else
        mbar_expect_tx(bar, row_bytes * (uint32_t)t.ny + (uint32_t)item.z * 16u);
#endif
        const float *src = sc.features + (((int64_t)t.bv * sc.H + t.y0) * sc.W + t.x0) * CN_FEAT_C;
        const uint32_t dst = smem_u32(smem) + slot * kRowsSlotBytes;
        bulk_g2s(dst + kTileBytes, g.rec_s + item.y, (uint32_t)item.z * 16u, bar);
#if CN_DIAG != 1
        for (int yy = 0; yy < t.ny; ++yy)
          bulk_g2s(dst + yy * (kTT * kTexBytes), src + (int64_t)yy * sc.W * CN_FEAT_C, row_bytes, bar);
#endif
        it = atomicAdd(cursor, 1);
        item = it < n_items ? g.items[it] : make_int4(0, 0, 0, 0);
#if CN_TILE_L2_PREFETCH
        // both slots are busy now (one being consumed, one landing): the next item's tile and records start their
        // way from DRAM to L2, so that the copy into the slot it will get is an L2 hit
        if (it < n_items) {
          const TileGeo tn = tile_geo(item.x, g.tiles_x, g.tiles_y, sc.H, sc.W);
          const float *srcn = sc.features + (((int64_t)tn.bv * sc.H + tn.y0) * sc.W + tn.x0) * CN_FEAT_C;
          for (int yy = 0; yy < tn.ny; ++yy)
            bulk_prefetch_l2(srcn + (int64_t)yy * sc.W * CN_FEAT_C, (uint32_t)tn.nx * kTexBytes);
          bulk_prefetch_l2(g.rec_s + item.y, (uint32_t)item.z * 16u);
        }
#endif
      }
    }
  } else {
    // =========================== consumers: one warp per row ===========================
    const uint32_t cw = warp - 1;
    if constexpr (!kBackward) {
      for (uint32_t k = 0;; ++k) {
        const uint32_t slot = k & 1u;
        mbar_wait(smem_u32(&bar_full[slot]), (k >> 1) & 1u);
        const int count = sdesc[slot].count;
        if (count == 0) break;
        const float *tile = reinterpret_cast<const float *>(smem + slot * kRowsSlotBytes);
        const int4 *recs = reinterpret_cast<const int4 *>(smem + slot * kRowsSlotBytes + kTileBytes);
        for (int j = cw; j < (CN_DIAG == 2 ? 0 : count); j += kRowsConsumerWarps) {   // CN_DIAG 2: loads only
          const int4 cur = recs[j];
          const float ax = __int_as_float(cur.z), ay = __int_as_float(cur.w);
          const Corners c = load_corners(tile, cur.y, l);
          uint8_t *xt = x_tiles + (int64_t)(cur.x >> 7) * CN_X_TILE_BYTES;
          const uint32_t rr = (uint32_t)cur.x & 127u;
          const float4 g0 = lerp2_f4(c.tl0, c.tr0, c.bl0, c.br0, ax, ay);
          const float4 g1 = lerp2_f4(c.tl1, c.tr1, c.bl1, c.br1, ax, ay);
          // channels 4l..4l+3 -> half of chunk l/2; channels 128+4l.. -> half of chunk 16 + l/2
          *reinterpret_cast<uint2 *>(xt + tile_chunk_off(rr, l >> 1) + (l & 1) * 8) =
              make_uint2(pack_bf16x2(g0.x, g0.y), pack_bf16x2(g0.z, g0.w));
          *reinterpret_cast<uint2 *>(xt + tile_chunk_off(rr, 16 + (l >> 1)) + (l & 1) * 8) =
              make_uint2(pack_bf16x2(g1.x, g1.y), pack_bf16x2(g1.z, g1.w));
        }
        __syncwarp();
        if (l == 0) mbar_arrive(smem_u32(&bar_empty[slot]));
      }
    } else {
      // A warp works in batches of 4 rows (rows j0 + 16 u of the item): the 8 dX loads of a batch are issued one
      // batch AHEAD of its math, across work items -- an item gives a warp only one or two batches, so a pipeline
      // inside the item would expose the full load latency once per item.  The first batch of item k+1 goes out
      // before the math of the warp's last batch of item k (its records are in the other slot, normally landed).
      constexpr int kStride = 4 * kRowsConsumerWarps;
      auto recs_of = [&](uint32_t kk) {
        return reinterpret_cast<const int4 *>(smem + (kk & 1u) * kRowsSlotBytes + kTileBytes);
      };
      auto issue = [&](const int4 *recs, int count, int j0, uint2 (&ga)[4], uint2 (&gb)[4]) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int4 r = recs[min(j0 + u * kRowsConsumerWarps, count - 1)];
          const uint8_t *t = x_tiles + (int64_t)(r.x >> 7) * CN_X_TILE_BYTES;
          const uint32_t rr = (uint32_t)r.x & 127u;
#if CN_DIAG == 3   // timing diagnostic: no dX loads
          ga[u] = make_uint2((uint32_t)(uintptr_t)t, rr);
          gb[u] = make_uint2(rr, (uint32_t)(uintptr_t)t);
#else
          ga[u] = __ldg(reinterpret_cast<const uint2 *>(t + tile_chunk_off(rr, l >> 1) + (l & 1) * 8));
          gb[u] = __ldg(reinterpret_cast<const uint2 *>(t + tile_chunk_off(rr, 16 + (l >> 1)) + (l & 1) * 8));
#endif
        }
      };
      uint32_t k = 0;
      mbar_wait(smem_u32(&bar_full[0]), 0);
      int count = sdesc[0].count;
      uint2 ga[4], gb[4];
      bool have = false;                 // ga / gb already hold the loads of the batch about to be computed
      while (count != 0) {
        const uint32_t slot = k & 1u;
        const float *tile = reinterpret_cast<const float *>(smem + slot * kRowsSlotBytes);
        const int4 *recs = recs_of(k);
        int next_count = -1;             // rows of item k+1, once looked at
        for (int j0 = cw; j0 < (CN_DIAG == 2 ? 0 : count); j0 += kStride) {
          if (!have) issue(recs, count, j0, ga, gb);
          uint2 na[4], nb[4];
          bool have_next = false;
          if (j0 + kStride < count) {
            issue(recs, count, j0 + kStride, na, nb);
            have_next = true;
          } else {
            const uint32_t nk = k + 1;
            mbar_wait(smem_u32(&bar_full[nk & 1u]), (nk >> 1) & 1u);
            next_count = sdesc[nk & 1u].count;
            if ((int)cw < next_count) {
              issue(recs_of(nk), next_count, (int)cw, na, nb);
              have_next = true;
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int j = j0 + u * kRowsConsumerWarps;
            if (j >= count) break;
            const int4 rec = recs[j];
            const float ax = __int_as_float(rec.z), ay = __int_as_float(rec.w);
            const Corners c = load_corners(tile, rec.y, l);
            const float gv[8] = {bf16_lo(ga[u].x), bf16_hi(ga[u].x), bf16_lo(ga[u].y), bf16_hi(ga[u].y),
                                 bf16_lo(gb[u].x), bf16_hi(gb[u].x), bf16_lo(gb[u].y), bf16_hi(gb[u].y)};
            const float tlv[8] = {c.tl0.x, c.tl0.y, c.tl0.z, c.tl0.w, c.tl1.x, c.tl1.y, c.tl1.z, c.tl1.w};
            const float trv[8] = {c.tr0.x, c.tr0.y, c.tr0.z, c.tr0.w, c.tr1.x, c.tr1.y, c.tr1.z, c.tr1.w};
            const float blv[8] = {c.bl0.x, c.bl0.y, c.bl0.z, c.bl0.w, c.bl1.x, c.bl1.y, c.bl1.z, c.bl1.w};
            const float brv[8] = {c.br0.x, c.br0.y, c.br0.z, c.br0.w, c.br1.x, c.br1.y, c.br1.z, c.br1.w};
            // four dot products <g, corner> (32 FMAs), combined per lane:
            //   d/d ax = (1-ay)(s_tr - s_tl) + ay (s_br - s_bl),  d/d ay = (1-ax)(s_bl - s_tl) + ax (s_br - s_tr)
            float s_tl = 0.f, s_tr = 0.f, s_bl = 0.f, s_br = 0.f;
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) {
              s_tl = fmaf(gv[ch], tlv[ch], s_tl);
              s_tr = fmaf(gv[ch], trv[ch], s_tr);
              s_bl = fmaf(gv[ch], blv[ch], s_bl);
              s_br = fmaf(gv[ch], brv[ch], s_br);
            }
            const float gax = (1.f - ay) * (s_tr - s_tl) + ay * (s_br - s_bl);
            const float gay = (1.f - ax) * (s_bl - s_tl) + ax * (s_br - s_tr);
            // one butterfly for both sums: after the first exchange the lower half-warp carries d/d ax, the upper d/d ay
            float v = (l < 16) ? gax : gay;
            v += __shfl_xor_sync(0xffffffffu, (l < 16) ? gay : gax, 16);
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            const float v_hi = __shfl_sync(0xffffffffu, v, 16);
            if (l == 0) g.gpix[rec.x] = make_float2(v, v_hi);
          }
          if (have_next) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              ga[u] = na[u];
              gb[u] = nb[u];
            }
          }
          have = have_next;
        }
        __syncwarp();
        if (l == 0) mbar_arrive(smem_u32(&bar_empty[slot]));
        ++k;
        if (next_count < 0) {
          mbar_wait(smem_u32(&bar_full[k & 1u]), (k >> 1) & 1u);
          next_count = sdesc[k & 1u].count;
        }
        count = next_count;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// feature-map gradient: every texel PULLS the rows of its (up to) four neighbouring cells
// ------------------------------------------------------------------------------------------
// The sorted order makes the rows of one (view, cell) a contiguous range [key_off[k], key_off[k+1]),
// so d_features[texel] = sum over the 4 cells whose 2x2 footprint contains the texel of
// weight(row, corner) * dX[row] is a segmented sum with no write conflicts: one warp owns a texel,
// accumulates its 256 channels in registers (lane l: channels 4l..4l+3 and 128+4l..128+4l+3) and
// writes the 1 KiB result once with plain coalesced stores -- no atomics, no zero-filled map, no
// read-modify-write traffic.  Several row sets (coarse + fine pass) are summed in the same sweep.
// Texels with more than kHeavyRows rows (border texels that collect every clamped out-of-frustum
// sample) are deferred to a second kernel in which a whole CTA splits the rows.
constexpr int kMaxSets = 4;
constexpr int kFgWarps = 8;                     // main sweep: 3 CTAs of 8 warps per SM (<= 85 registers)
constexpr int kFgThreads = 32 * kFgWarps;
constexpr int kFgCtasPerSm = 3;
constexpr int kHeavyRows = 192;                 // a texel with more rows is deferred to the second kernel ...
constexpr int kHeavyChunk = 256;                // ... as independent chunks of this many rows
constexpr int kUnitRows = 2;  // a work unit = kUnitRows x 8 texels of one 8x8 texel tile

struct GradSet {
  const int *key_off;    // [n_keys] start of every key in rec_s
  const int4 *rec_s;     // sorted records
  const uint8_t *dx;     // dX tile images of this pass
  int rows;
};
struct GradArgs {
  GradSet set[kMaxSets];
  int n_sets;
  int n_keys, tiles_x, tiles_y;  // cell tiling of the records (same scene for every set)
  int ttx, tty;                  // texel tiling: ceil(W/8), ceil(H/8)
  int n_units;
  int accumulate;
  int *cursor;      // work-unit cursor
  int *heavy_n;     // number of deferred entries, and the cursor of the second kernel
  int2 *heavy_list; // [heavy_cap] (texel index (bv*H + Y)*W + X, first flat row of the chunk)
  int heavy_cap;
};

// (range start, length) of the rows of pair p = set*4 + corner for texel (bv, Y, X); corner = dy*2 + dx
__device__ __forceinline__ void texel_pair_range(const CnScene &sc, const GradArgs &a, int p, int bv, int Y, int X,
                                                 int &st, int &len) {
  st = 0;
  len = 0;
  if (p >= a.n_sets * 4) return;
  const int fy = Y - ((p >> 1) & 1), fx = X - (p & 1);
  if (fy < 0 || fy > sc.H - 2 || fx < 0 || fx > sc.W - 2) return;
  const int key = ((bv * a.tiles_y + fy / kT) * a.tiles_x + fx / kT) * kCells + (fy % kT) * kT + (fx % kT);
  const GradSet &gs = a.set[p >> 2];
  st = __ldg(gs.key_off + key);
  const int en = key + 1 < a.n_keys ? __ldg(gs.key_off + key + 1) : gs.rows;
  len = en - st;
}

// lane f of the warp: record of flat row b0 + f of this texel's pairs (raw, not yet consumed), and its pair
struct RowFetch {
  int4 rec;
  int pair;   // set * 4 + corner, -1 = no row
};
__device__ __forceinline__ RowFetch fetch_rows(const GradArgs &a, int st, int len, int off_incl, int n, int b0) {
  const uint32_t l = lane_id();
  const int n_pairs = a.n_sets * 4;
  const int f = b0 + (int)l;
  int my_p = 0;
  for (int p = 0; p < n_pairs; ++p) my_p += (f >= __shfl_sync(0xffffffffu, off_incl, p)) ? 1 : 0;
  my_p = min(my_p, n_pairs - 1);
  const int p_st = __shfl_sync(0xffffffffu, st, my_p);
  const int p_off = __shfl_sync(0xffffffffu, off_incl - len, my_p);
  RowFetch r;
  r.pair = -1;
  r.rec = make_int4(0, 0, 0, 0);
  if (f < n) {
    r.pair = my_p;
    r.rec = __ldg(a.set[my_p >> 2].rec_s + p_st + (f - p_off));
  }
  return r;
}

// rows u0 .. u0+U-1 (masked at nb) of the batch held by the warp's lanes: all 2U loads first, then the FMAs
template <int U>
__device__ __forceinline__ void accumulate_group(float w, uint32_t a_lo, uint32_t a_hi, uint32_t lane_off, int u0,
                                                 int nb, float (&acc)[8]) {
  uint2 va[U], vb[U];
  float wu[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int src = min(u0 + u, nb - 1);
    wu[u] = (u0 + u < nb) ? __shfl_sync(0xffffffffu, w, src) : 0.f;
    const uint32_t lo = __shfl_sync(0xffffffffu, a_lo, src), hi = __shfl_sync(0xffffffffu, a_hi, src);
    const uint64_t base = ((uint64_t)hi << 32) | (lo & ~127u);
    const uint8_t *p = reinterpret_cast<const uint8_t *>(base + (lane_off ^ ((lo & 7u) << 4)));
    va[u] = __ldg(reinterpret_cast<const uint2 *>(p));
    vb[u] = __ldg(reinterpret_cast<const uint2 *>(p + 2 * CN_SLAB_BYTES));
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    acc[0] = fmaf(wu[u], bf16_lo(va[u].x), acc[0]); acc[1] = fmaf(wu[u], bf16_hi(va[u].x), acc[1]);
    acc[2] = fmaf(wu[u], bf16_lo(va[u].y), acc[2]); acc[3] = fmaf(wu[u], bf16_hi(va[u].y), acc[3]);
    acc[4] = fmaf(wu[u], bf16_lo(vb[u].x), acc[4]); acc[5] = fmaf(wu[u], bf16_hi(vb[u].x), acc[5]);
    acc[6] = fmaf(wu[u], bf16_lo(vb[u].y), acc[6]); acc[7] = fmaf(wu[u], bf16_hi(vb[u].y), acc[7]);
  }
}

// acc += sum over the nb rows held by lanes 0..nb-1 of  weight(row, corner) * dX[row]
__device__ __forceinline__ void accumulate_rows(const GradArgs &a, const RowFetch &rf, int nb, float (&acc)[8]) {
  const uint32_t l = lane_id();
  // lane constant part of the dX address: channels 4l..4l+3 live in chunk l/2 (half l%2) of slabs 0..1,
  // channels 128+4l.. in the same position two slabs further
  const uint32_t lane_off = ((l >> 4) * CN_SLAB_BYTES) + (((l >> 1) & 7u) << 4) + (l & 1u) * 8u;
  float w = 0.f;
  uint64_t addr = 0;
  if (rf.pair >= 0) {
    const float ax = __int_as_float(rf.rec.z), ay = __int_as_float(rf.rec.w);
    w = ((rf.pair & 1) ? ax : 1.f - ax) * ((rf.pair & 2) ? ay : 1.f - ay);
    const uint32_t row = (uint32_t)rf.rec.x;
    // 128-byte aligned row start inside the dX image; the swizzle phase (row % 8) rides in the low bits
    addr = (uint64_t)(uintptr_t)a.set[rf.pair >> 2].dx +
           ((uint64_t)(row >> 7) * CN_X_TILE_BYTES + (row & 127u) * 128u) + (row & 7u);
  }
  const uint32_t a_lo = (uint32_t)addr, a_hi = (uint32_t)(addr >> 32);
  // groups of 8 rows (16 loads in flight per lane) while more than 4 rows remain, else a group of 4
  int u0 = 0;
  while (u0 < nb) {
    if (nb - u0 > 4) {
      accumulate_group<8>(w, a_lo, a_hi, lane_off, u0, nb, acc);
      u0 += 8;
    } else {
      accumulate_group<4>(w, a_lo, a_hi, lane_off, u0, nb, acc);
      u0 += 4;
    }
  }
}

// inclusive scan of `len` over the first 16 lanes
__device__ __forceinline__ int pair_scan(int len, int &total) {
  const uint32_t l = lane_id();
  int inc = len;
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, inc, o);
    if (l >= (uint32_t)o) inc += v;
  }
  total = __shfl_sync(0xffffffffu, inc, 15);
  return inc;
}

__device__ __forceinline__ void texel_store(float *__restrict__ d_features, int64_t texel, const float (&acc)[8],
                                            int accumulate) {
  const uint32_t l = lane_id();
  float4 *o = reinterpret_cast<float4 *>(d_features + texel * CN_FEAT_C);
  float4 r0 = make_float4(acc[0], acc[1], acc[2], acc[3]), r1 = make_float4(acc[4], acc[5], acc[6], acc[7]);
  if (accumulate) {
    const float4 p0 = o[l], p1 = o[32 + l];
    r0.x += p0.x; r0.y += p0.y; r0.z += p0.z; r0.w += p0.w;
    r1.x += p1.x; r1.y += p1.y; r1.z += p1.z; r1.w += p1.w;
  }
  o[l] = r0;
  o[32 + l] = r1;
}

// Per warp, three dependent global loads lead to a texel's sum (key offsets -> records -> dX rows), so
// the texels of a work unit run through a 3-stage software pipeline: while the dX rows of texel t
// are summed, the records of texel t+1 and the key offsets of texel t+2 are already in flight.
__global__ void __launch_bounds__(kFgThreads, kFgCtasPerSm)
feature_grad_kernel(CnScene sc, GradArgs a, float *__restrict__ d_features) {
  const uint32_t l = lane_id();
  constexpr int kUnitTexels = kUnitRows * kT;
  int next_unit = 0;
  if (l == 0) next_unit = atomicAdd(a.cursor, 1);
  for (;;) {
    const int unit = __shfl_sync(0xffffffffu, next_unit, 0);
    if (unit >= a.n_units) break;
    if (l == 0) next_unit = atomicAdd(a.cursor, 1);  // consumed at the top of the next round
    const int q = unit % (kT / kUnitRows);
    const int tile = unit / (kT / kUnitRows);
    const int tx = tile % a.ttx;
    const int ty = (tile / a.ttx) % a.tty;
    const int bv = tile / (a.ttx * a.tty);
    const int Y0 = ty * kT + q * kUnitRows, X0 = tx * kT;
    auto ranges = [&](int t, int &st, int &len) {  // stage A
      st = 0;
      len = 0;
      const int Y = Y0 + t % kUnitRows, X = X0 + t / kUnitRows;  // column-major: the 4 texels of a cell are adjacent in time
      if (t < kUnitTexels && Y < sc.H && X < sc.W) texel_pair_range(sc, a, (int)l, bv, Y, X, st, len);
    };
    int st1, len1, st2, len2;          // ranges of texel t+1 (loaded), t+2 (in flight)
    int st0, len0, off0, n0;           // texel t: ranges, scan
    RowFetch rf0;                      // texel t: first 32 records (in flight)
    ranges(0, st0, len0);
    ranges(1, st1, len1);
    off0 = pair_scan(len0, n0);
    rf0 = fetch_rows(a, st0, len0, off0, min(n0, n0 > kHeavyRows ? 0 : 32), 0);
    for (int t = 0; t < kUnitTexels; ++t) {
      ranges(t + 2, st2, len2);
      // stage B for texel t+1
      int n1;
      const int off1 = pair_scan(len1, n1);
      const RowFetch rf1 = fetch_rows(a, st1, len1, off1, min(n1, n1 > kHeavyRows ? 0 : 32), 0);
      // stage C for texel t
      const int Y = Y0 + t % kUnitRows, X = X0 + t / kUnitRows;
      if (Y < sc.H && X < sc.W) {
        const int64_t texel = ((int64_t)bv * sc.H + Y) * sc.W + X;
        bool deferred = false;
        if (n0 > kHeavyRows) {
          const int n_chunks = (n0 + kHeavyChunk - 1) / kHeavyChunk;
          int slot = 0;
          if (l == 0) slot = atomicAdd(a.heavy_n, n_chunks);
          slot = __shfl_sync(0xffffffffu, slot, 0);
          if (slot + n_chunks <= a.heavy_cap) {
            for (int c = (int)l; c < n_chunks; c += 32) a.heavy_list[slot + c] = make_int2((int)texel, c * kHeavyChunk);
            deferred = true;
            if (!a.accumulate) {  // the chunks are summed with red.global.add: start from zero
              const float z8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
              texel_store(d_features, texel, z8, 0);
            }
          } else {
            // list full (unreachable with the list sized by heavy_cap_for; kept correct): the slots this texel
            // reserved below the cap are published as skip entries -- heavy_n has already moved past them and the
            // second kernel walks [0, min(heavy_n, cap)) -- and the texel is summed here after all
            for (int c = slot + (int)l; c < min(slot + n_chunks, a.heavy_cap); c += 32) a.heavy_list[c] = make_int2(-1, 0);
            rf0 = fetch_rows(a, st0, len0, off0, n0, 0);
          }
        }
        if (!deferred) {
          float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          accumulate_rows(a, rf0, min(32, n0), acc);
          for (int b0 = 32; b0 < n0; b0 += 32) {
            const RowFetch rf = fetch_rows(a, st0, len0, off0, n0, b0);
            accumulate_rows(a, rf, min(32, n0 - b0), acc);
          }
          texel_store(d_features, texel, acc, a.accumulate);
        }
      }
      st0 = st1; len0 = len1; off0 = off1; n0 = n1; rf0 = rf1;
      st1 = st2; len1 = len2;
    }
  }
}

// deferred texels (frame borders collect every clamped out-of-frustum sample): independent chunks of
// kHeavyChunk rows, one warp each, partial sums added with red.global.add.v4.f32
__global__ void __launch_bounds__(kFgThreads, kFgCtasPerSm)
feature_grad_heavy_kernel(CnScene sc, GradArgs a, float *__restrict__ d_features) {
  const uint32_t l = lane_id();
  const int n_entries = min(a.heavy_n[0], a.heavy_cap);
  for (;;) {
    int e = 0;
    if (l == 0) e = atomicAdd(a.heavy_n + 1, 1);
    e = __shfl_sync(0xffffffffu, e, 0);
    if (e >= n_entries) break;
    const int2 ent = a.heavy_list[e];
    if (ent.x < 0) continue;                       // skip entry of a texel that did not fit (summed by the main sweep)
    const int64_t texel = ent.x;
    const int X = (int)(texel % sc.W);
    const int Y = (int)((texel / sc.W) % sc.H);
    const int bv = (int)(texel / ((int64_t)sc.W * sc.H));
    int st, len, n;
    texel_pair_range(sc, a, (int)l, bv, Y, X, st, len);
    const int off_incl = pair_scan(len, n);
    const int end = min(n, ent.y + kHeavyChunk);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    RowFetch rf = fetch_rows(a, st, len, off_incl, end, ent.y);
    for (int b0 = ent.y; b0 < end; b0 += 32) {
      RowFetch rf_next;
      rf_next.pair = -1;
      rf_next.rec = make_int4(0, 0, 0, 0);
      if (b0 + 32 < end) rf_next = fetch_rows(a, st, len, off_incl, end, b0 + 32);  // in flight during the sums
      accumulate_rows(a, rf, min(32, end - b0), acc);
      rf = rf_next;
    }
    float *o = d_features + texel * CN_FEAT_C;
    red_add_f4(o + 4 * l, acc[0], acc[1], acc[2], acc[3]);
    red_add_f4(o + 128 + 4 * l, acc[4], acc[5], acc[6], acc[7]);
  }
}

// ------------------------------------------------------------------------------------------
// d/dz: posenc'(p), image and pixel-coordinate gradients chained through K and E^-1
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
project_dz_kernel(CnScene sc, const float *__restrict__ ray_o, const float *__restrict__ ray_d,
                  const float *__restrict__ z, int nr, int S, const uint8_t *__restrict__ dx_tiles,
                  const float2 *__restrict__ gpix, float *__restrict__ dz, float *__restrict__ d_ray_o,
                  float *__restrict__ d_ray_d) {
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
  // p = E [o + z d; 1]  ->  d/dx_j = sum_i dp_i E_ij,  d/dz = sum_j (d/dx_j) d_j
  float gx[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) gx[j] = dp[0] * E[j] + dp[1] * E[4 + j] + dp[2] * E[8 + j];
  if (dz) atomicAdd(dz + id.n, gx[0] * rc.d[0] + gx[1] * rc.d[1] + gx[2] * rc.d[2]);
  if (d_ray_o) {
    // pose-query models (delta_ngf/model.py:196-232): the query point is the ray origin and its approach
    // direction the ray direction; d/d(origin) = d/dx, d/d(direction) = z d/dx + E3^T d/d(dcam), the latter
    // from the posenc(direction) columns of dX (slab 4 words 30,31 and slab 5 words 0..27)
    const float zz = z[id.n];
    float dd[3] = {0.f, 0.f, 0.f};
    float dcam[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) dcam[c] = dot4_rn(E + 4 * c, rc.d[0], rc.d[1], rc.d[2], 1.f);
    const uint4 v7 = *reinterpret_cast<const uint4 *>(my4 + ((7u ^ sw) << 4));
    uint32_t w[30];
    w[0] = v7.z;
    w[1] = v7.w;
#pragma unroll
    for (int q = 0; q < 7; ++q) {
      const uint4 v = *reinterpret_cast<const uint4 *>(my5 + ((q ^ sw) << 4));
      w[2 + 4 * q] = v.x; w[3 + 4 * q] = v.y; w[4 + 4 * q] = v.z; w[5 + 4 * q] = v.w;
    }
#pragma unroll
    for (int l = 0; l < 30; ++l) {
      const int c = l / 10, k = l % 10;
      const float fk = pe_freq(k);
      float s, co;
      sincos_posenc(__fmul_rn(dcam[c], fk), s, co);
      dd[c] += fk * (bf16_lo(w[l]) * co - bf16_hi(w[l]) * s);
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      atomicAdd(d_ray_o + id.ray * 3 + j, gx[j]);
      atomicAdd(d_ray_d + id.ray * 3 + j, zz * gx[j] + dd[0] * E[j] + dd[1] * E[4 + j] + dd[2] * E[8 + j]);
    }
  }
}

}  // namespace cn

#include "gather_lowres.cuh"

namespace cn {

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
  *out = ws_layout(scene->b, scene->nv, scene->H, scene->W, scene->feat_up, rows, ws);
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
  const int n_slots = 127 / S + 2;  // rays that can have rows in one 128-row tile
  project_bin_kernel<<<(unsigned)(g.rows / 128), 128, (size_t)n_slots * 30 * 4, st>>>(*scene, ray_o, ray_d, z, nr, S, g,
                                                                                   x_tiles);
  bin_scan1_kernel<<<g.n_blk, 256, 0, st>>>(g);
  bin_scan2_kernel<<<g.n_blk, 256, 0, st>>>(g);
  bin_fill_kernel<<<(unsigned)((g.rows + 255) / 256), 256, 0, st>>>(g);
  return check_launch("cn_project_bin");
}

template <bool kBackward>
static int run_tile_rows(const CnScene *scene, const Ws &g, int cursor_index, uint8_t *x_tiles, cudaStream_t st) {
  if (scene->feat_up == 2) {
    cudaError_t e = cudaFuncSetAttribute(tile_rows_lr_kernel<kBackward>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kRowsSmemLr);
    if (e != cudaSuccess) {
      set_error("tile_rows_lr: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return (int)e;
    }
    tile_rows_lr_kernel<kBackward><<<num_sms(), lr_threads<kBackward>(), kRowsSmemLr, st>>>(*scene, g, g.hdr + cursor_index,
                                                                                            x_tiles);
    return check_launch("tile_rows_lr");
  }
  cudaError_t e = cudaFuncSetAttribute(tile_rows_kernel<kBackward>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kRowsSmem);
  if (e != cudaSuccess) {
    set_error("tile_rows: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)e;
  }
  tile_rows_kernel<kBackward><<<num_sms(), rows_threads<kBackward>(), kRowsSmem, st>>>(*scene, g, g.hdr + cursor_index, x_tiles);
  return check_launch("tile_rows");
}


static int run_feature_grad_lr(const CnScene *scene, int n_sets, const Ws *const *ws, const void *const *dx,
                               float *d_features, int accumulate, cudaStream_t st) {
  CN_REQUIRE(n_sets <= kLrMaxSets, "cn_feature_grad: feat_up == 2 takes at most %d row sets (got %d)", kLrMaxSets, n_sets);
  LrGradArgs a;
  a.n_sets = n_sets;
  for (int i = 0; i < kLrMaxSets; ++i) {
    const Ws *w = ws[i < n_sets ? i : 0];
    a.set[i].key_off = w->count;
    a.set[i].rec_s = w->rec_s;
    a.set[i].dx = (const uint8_t *)dx[i < n_sets ? i : 0];
    a.set[i].rows = (int)w->rows;
  }
  const Ws *w0 = ws[0];
  a.n_keys = w0->n_keys;
  a.tiles_x = w0->tiles_x;
  a.tiles_y = w0->tiles_y;
  CN_REQUIRE((int64_t)w0->n_bins * kT * 4 < ((int64_t)1 << 31), "cn_feature_grad: too many cell rows");
  a.n_units = w0->n_bins * kT;
  a.cursor = w0->hdr + kHdrCurSc;
  const Ws *wl = w0;
  for (int i = 1; i < n_sets; ++i)
    if (ws[i]->rows > wl->rows) wl = ws[i];
  a.heavy_n = w0->hdr + kHdrHeavyN;
  a.heavy_list = wl->heavy;
  a.heavy_cap = wl->heavy_cap;
  if (const char *dbg = getenv("CN_DEBUG_HEAVY_CAP")) {   // tests only: force the list-full path
    const int cap = atoi(dbg);
    if (cap >= 0 && cap < a.heavy_cap) a.heavy_cap = cap;
  }
  cudaError_t ce = cudaMemsetAsync(w0->hdr + kHdrCurSc, 0, 3 * sizeof(int), st);
  if (ce == cudaSuccess && !accumulate)   // every contribution arrives as a reduction
    ce = cudaMemsetAsync(d_features, 0, (size_t)scene->b * scene->nv * (scene->H / 2) * (scene->W / 2) * CN_FEAT_C * 4, st);
  if (ce != cudaSuccess) {
    set_error("cn_feature_grad: cudaMemsetAsync: %s", cudaGetErrorString(ce));
    return (int)ce;
  }
  feature_grad_lr_kernel<<<kLrCtasPerSm * num_sms(), kLrThreads, 0, st>>>(*scene, a, d_features);
  feature_grad_lr_heavy_kernel<<<kLrCtasPerSm * num_sms(), kLrThreads, 0, st>>>(*scene, a, d_features);
  return check_launch("cn_feature_grad (feat_up 2)");
}

static int run_feature_grad(const CnScene *scene, int n_sets, const Ws *const *ws, const void *const *dx,
                            float *d_features, int accumulate, cudaStream_t st) {
  if (scene->feat_up == 2) return run_feature_grad_lr(scene, n_sets, ws, dx, d_features, accumulate, st);
  GradArgs a;
  a.n_sets = n_sets;
  for (int i = 0; i < kMaxSets; ++i) {
    const Ws *w = ws[i < n_sets ? i : 0];
    a.set[i].key_off = w->count;
    a.set[i].rec_s = w->rec_s;
    a.set[i].dx = (const uint8_t *)dx[i < n_sets ? i : 0];
    a.set[i].rows = (int)w->rows;
  }
  const Ws *w0 = ws[0];
  a.n_keys = w0->n_keys;
  a.tiles_x = w0->tiles_x;
  a.tiles_y = w0->tiles_y;
  a.ttx = (scene->W + kT - 1) / kT;
  a.tty = (scene->H + kT - 1) / kT;
  a.n_units = scene->b * scene->nv * a.ttx * a.tty * (kT / kUnitRows);
  a.accumulate = accumulate;
  a.cursor = w0->hdr + kHdrCurSc;
  // the deferred-chunk list of the largest set holds the chunks of every set (see heavy_cap_for)
  const Ws *wl = w0;
  for (int i = 1; i < n_sets; ++i)
    if (ws[i]->rows > wl->rows) wl = ws[i];
  a.heavy_n = w0->hdr + kHdrHeavyN;
  a.heavy_list = wl->heavy;
  a.heavy_cap = wl->heavy_cap;
  if (const char *dbg = getenv("CN_DEBUG_HEAVY_CAP")) {   // tests only: force the list-full path
    const int cap = atoi(dbg);
    if (cap >= 0 && cap < a.heavy_cap) a.heavy_cap = cap;
  }
  cudaError_t ce = cudaMemsetAsync(w0->hdr + kHdrCurSc, 0, 3 * sizeof(int), st);  // cursor + deferred entries + their cursor
  if (ce != cudaSuccess) {
    set_error("cn_feature_grad: cudaMemsetAsync: %s", cudaGetErrorString(ce));
    return (int)ce;
  }
  feature_grad_kernel<<<kFgCtasPerSm * num_sms(), kFgThreads, 0, st>>>(*scene, a, d_features);
  feature_grad_heavy_kernel<<<kFgCtasPerSm * num_sms(), kFgThreads, 0, st>>>(*scene, a, d_features);
  return check_launch("cn_feature_grad");
}

}  // namespace cn

extern "C" size_t cn_gather_ws_bytes(int32_t b, int32_t nv, int32_t H, int32_t W, int32_t nr, int32_t S) {
  if (b <= 0 || nv <= 0 || H < 2 || W < 2 || nr <= 0 || S <= 0) return 0;
  return cn::ws_layout(b, nv, H, W, 1, (int64_t)b * nv * nr * S, nullptr).total;
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
  cudaError_t ce = cudaMemsetAsync(g.hdr + kHdrCurZg, 0, sizeof(int), st);
  if (ce != cudaSuccess) {
    set_error("cn_project_gather_bwd: cudaMemsetAsync: %s", cudaGetErrorString(ce));
    return (int)ce;
  }
  if (dz) {
    if (int e = run_tile_rows<true>(scene, g, kHdrCurZg, (uint8_t *)dx_tiles, st)) return e;
    project_dz_kernel<<<(unsigned)(g.rows / 128), 128, 0, st>>>(*scene, ray_o, ray_d, z, nr, S,
                                                               (const uint8_t *)dx_tiles, g.gpix, dz, nullptr, nullptr);
    if (int e = check_launch("cn_project_gather_bwd (dz)")) return e;
  }
  if (d_features) {
    const Ws *gp = &g;
    const void *dxp = dx_tiles;
    if (int e = run_feature_grad(scene, 1, &gp, &dxp, d_features, 1, st)) return e;
  }
  return 0;
}

extern "C" int cn_project_gather_bwd_rays(const CnScene *scene, const float *ray_o, const float *ray_d, const float *z,
                                          int32_t nr, int32_t S, const void *dx_tiles, void *ws, size_t ws_bytes,
                                          float *dz, float *d_ray_o, float *d_ray_d, void *stream) {
  using namespace cn;
  if (int e = check_scene(scene, "cn_project_gather_bwd_rays")) return e;
  CN_REQUIRE(ray_o && ray_d && z && dx_tiles && d_ray_o && d_ray_d, "cn_project_gather_bwd_rays: null pointer");
  CN_REQUIRE(((uintptr_t)dx_tiles & 127) == 0, "cn_project_gather_bwd_rays: dx_tiles must be 128-byte aligned");
  Ws g;
  if (int e = get_ws(scene, nr, S, ws, ws_bytes, "cn_project_gather_bwd_rays", &g)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t ce = cudaMemsetAsync(g.hdr + kHdrCurZg, 0, sizeof(int), st);
  if (ce != cudaSuccess) {
    set_error("cn_project_gather_bwd_rays: cudaMemsetAsync: %s", cudaGetErrorString(ce));
    return (int)ce;
  }
  if (int e = run_tile_rows<true>(scene, g, kHdrCurZg, (uint8_t *)dx_tiles, st)) return e;
  project_dz_kernel<<<(unsigned)(g.rows / 128), 128, 0, st>>>(*scene, ray_o, ray_d, z, nr, S, (const uint8_t *)dx_tiles,
                                                             g.gpix, dz, d_ray_o, d_ray_d);
  return check_launch("cn_project_gather_bwd_rays");
}

extern "C" int cn_feature_grad(const CnScene *scene, int32_t n_sets, const int32_t *nr, const int32_t *S,
                               void *const *ws, const size_t *ws_bytes, const void *const *dx_tiles,
                               float *d_features, int32_t accumulate, void *stream) {
  using namespace cn;
  if (int e = check_scene(scene, "cn_feature_grad")) return e;
  CN_REQUIRE(n_sets >= 1 && n_sets <= kMaxSets, "cn_feature_grad: n_sets must be 1..%d (got %d)", kMaxSets, n_sets);
  CN_REQUIRE(nr && S && ws && ws_bytes && dx_tiles && d_features, "cn_feature_grad: null pointer");
  CN_REQUIRE(((uintptr_t)d_features & 15) == 0, "cn_feature_grad: d_features alignment");
  CN_REQUIRE((int64_t)scene->b * scene->nv * scene->H * scene->W < ((int64_t)1 << 31), "cn_feature_grad: map too large");
  Ws g[kMaxSets];
  const Ws *gp[kMaxSets];
  for (int i = 0; i < n_sets; ++i) {
    CN_REQUIRE(dx_tiles[i] && ((uintptr_t)dx_tiles[i] & 127) == 0, "cn_feature_grad: dx_tiles[%d] null or misaligned", i);
    if (int e = get_ws(scene, nr[i], S[i], ws[i], ws_bytes[i], "cn_feature_grad", &g[i])) return e;
    gp[i] = &g[i];
  }
  return run_feature_grad(scene, n_sets, gp, dx_tiles, d_features, accumulate, (cudaStream_t)stream);
}
