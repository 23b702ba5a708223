/*
 * clipnerf_b200 -- C ABI of the B200 (sm_100a) CLIP-NeRF volumetric render path.
 *
 * This is the drop-in boundary of the repo: every entry point replaces one piece of
 * the reference's TensorFlow hot path (TWeber132/thesis-clip-nerf; citations are
 * file:line in that repository).  The reference has no FFI of its own (it is pure
 * TensorFlow/Keras), so these are the functions a maintainer would bind from the
 * reference's Python modules (see INTEGRATION.md for the ctypes stubs).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no C++ / torch types.
 *   - every pointer is a DEVICE pointer unless the name ends in _host.
 *   - the caller allocates every output and workspace; the library owns no device
 *     memory and keeps no global mutable state except a thread-local error string.
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*); nothing here
 *     synchronises with the host.
 *   - return value: 0 = ok, < 0 = argument / shape / alignment error
 *     (text via cn_last_error()), > 0 = a cudaError_t.
 *   - there is no CPU fallback.
 *
 * Row order ("sample-view rows"): the per-view MLP stage works on rows
 *   row = ((b * nv + v) * nr + r) * S + i      (b-major "(b nv) nr np", layers.py:366)
 * and the fusion stage on rows  q = (b * nr + r) * S + i.  Rows are grouped in
 * tiles of CN_TILE_ROWS; nr * S must be a multiple of CN_TILE_ROWS.
 *
 * bf16 tile images: a [128 rows x 64 cols] bf16 "slab" is 16 KiB, row r at byte
 * r*128, the 16-byte chunk c of a row stored at chunk position c ^ (r & 7)
 * (the 128-byte swizzle tcgen05 shared-memory descriptors expect).  An activation
 * tile [128 x 128] is 2 slabs (32 KiB), an MLP-input tile [128 x 384] is 6 slabs.
 */
#ifndef CLIPNERF_B200_H
#define CLIPNERF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CN_ABI_VERSION 11
#define CN_TILE_ROWS 128
#define CN_HIDDEN 128
#define CN_FEAT_C 256          /* combined_features channels (model_v0.py:60)            */
#define CN_IN_DIM 379          /* 60 + 60 + 3 + 256 (layers.py:362)                      */
#define CN_IN_PAD 384          /* K of the first Dense, padded to a multiple of 64       */
#define CN_SLAB_BYTES 16384
#define CN_ACT_TILE_BYTES 32768
#define CN_X_TILE_BYTES 98304

/* number of fp32 parameters of one MVResNetMLPNeRFEmbedding(6,128) + RenderReadout(4)
 * in the flat layout described at cn_pack_weights(). */
#define CN_EMB_PARAMS 246784
#define CN_READOUT_PARAMS 516
#define CN_NET_PARAMS (CN_EMB_PARAMS + CN_READOUT_PARAMS)

/* bytes of the packed (bf16 operand images + fp32 bias tables) form of one net */
#define CN_PACKED_BYTES (2 * 491520 + 16384)

/* GraspReadout (delta_ngf/layers.py:8-28): fp32 parameters in Keras variable order, kernel[in][out] --
 * activation_downscale_{1..4} (kernel [128][64], bias [64]), combined_activation_downscale ([256][64], [64]),
 * readout.0 layer_0 ([2688][128], [128]), layer_1 ([128][64], [64]), shortcut ([2688][64]), readout.1 layer_0 / layer_1
 * ([64][64], [64] each), readout.2 output_layer ([64][1], [1]; zero when use_bias is False) -- and the size of their
 * packed form (bf16 operand images in both orientations + fp32 bias table, cn_grasp_head_pack). */
#define CN_HEAD_PARAMS 582337
#define CN_HEAD_PACK_BYTES 2329600
#define CN_HEAD_OFFSETS 42     /* query offsets per pose: 6 gripper bases x n_5d_poses = 7 (delta_ngf/model.py:63-97) */

int cn_abi_version(void);
const char *cn_last_error(void);

/* Source views of one batch of scenes: the 5-tuple of MVNeRFDataGenerator.get_input
 * (src/lib/data_generator/mvnerf.py:28-43) plus combined_features (model_v0.py:60). */
typedef struct CnScene {
  const float *images;     /* [b, nv, H, W, 3]  in [0,1]                  */
  const float *features;   /* [b, nv, H / feat_up, W / feat_up, 256]      */
  const float *intrinsics; /* [b, nv, 4, 4]                               */
  const float *extr_inv;   /* [b, nv, 4, 4]  world -> camera              */
  int32_t b, nv, H, W;
  /* 1: `features` is combined_features itself (model_v0.py:60).
   * 2: `features` is the INPUT of the encoder's final UpSampling2D(2, 'bilinear') (layers.py:589,659;
   *    legacy_layers.py:190; model_without.py:76): every gather evaluates
   *    interpolate_bilinear(UpSampling2D(x)) as the composed <= 3x3 stencil on the [H/2, W/2] map (fixed
   *    0.25 / 0.75 weights, half-pixel centres, edge clamp) and every d_features is the gradient w.r.t. that
   *    low-resolution map; H and W must be even.  Pixel coordinates, images and dz are unchanged.          */
  int32_t feat_up;
} CnScene;

/* ---- sampling ------------------------------------------------------------------ */

/* sample_along_ray (nerf_utils.py:49-61): z[n, S] = fp32(near + i*step) + u[n, S]*step */
int cn_sample_coarse(const float *u, float *z, int64_t n_rays, int32_t S, double near_, double far_,
                     void *stream);

/* sample_pdf (nerf_utils.py:143-176) + concat + sort (model_v0.py:150-156), S = 64.
 *   z_c, w_c, u : [n, 64]   ->  z_fine [n, 128] ascending (stable),
 *   perm  [n,128] u8 : source index of each sorted value (0..63 coarse, 64..127 new),
 *   above [n, 64] u8 : inverse-CDF bin index of each draw (1..63).                      */
int cn_sample_pdf_fwd(const float *z_c, const float *w_c, const float *u, int64_t n_rays,
                      float *z_fine, uint8_t *perm, uint8_t *above, void *stream);
/* op-level sample_pdf with the reference's own signature (nerf_utils.py:143): bins [n,63],
 * weights [n,62], u [n,64] -> samples [n,64] (unsorted), above [n,64] int32 (nullable).        */
int cn_sample_pdf(const float *bins, const float *weights, const float *u, int64_t n_rays,
                  float *samples, int32_t *above, void *stream);
/* gradient of the above w.r.t. the coarse weights (no stop_gradient in the reference,
 * model_v0.py:153-158):  dz_fine [n,128] -> dw_c [n,64] (accumulated, +=).              */
int cn_sample_pdf_bwd(const float *z_c, const float *w_c, const float *u, const uint8_t *perm,
                      const float *dz_fine, int64_t n_rays, float *dw_c, void *stream);

/* ---- projection + bilinear gather + positional encoding ------------------------- */

/* The gather and its backward work on SPATIALLY BINNED sample-view rows: every row is keyed by
 * the 8x8-texel tile (and the cell inside it) that its clamped bilinear footprint falls in, the
 * rows are counting-sorted by key, and each CTA stages one 9x9-texel tile of the feature map in
 * shared memory (cp.async.bulk) and serves all rows of that tile from there; the backward
 * accumulates the tile's gradient on chip and adds it to d_features with one bulk reduction
 * (cp.reduce.async.bulk) per tile row.  The sorted row records live in a caller-allocated
 * workspace that the forward fills and the backward re-uses.                                   */
size_t cn_gather_ws_bytes(int32_t b, int32_t nv, int32_t H, int32_t W, int32_t nr, int32_t S);

/* projection + binning only (fills `ws`; no MLP-input rows are written)                        */
int cn_project_bin(const CnScene *scene, const float *ray_o, const float *ray_d, const float *z,
                   int32_t nr, int32_t S, void *ws, size_t ws_bytes, void *stream);

/* compute_pixel_in_image_mv + get_projection_features_mv +
 * world_to_camera_direction_vector_mv + position_encoding + the concat of
 * layers.py:362 (nerf_utils.py:64-126,277-285): writes the MLP input rows as bf16
 * tile images X [b*nv*nr*S/128 tiles][6 slabs].  Column order inside X:
 * [feat 0..255 | posenc(p_cam) 60 | posenc(d_cam) 60 | 2*img-1 (3) | 0 (5)].
 * Fills `ws` (cn_gather_ws_bytes) with the binned row records.                                 */
int cn_project_gather_fwd(const CnScene *scene, const float *ray_o, const float *ray_d,
                          const float *z, int32_t nr, int32_t S, void *x_tiles, void *ws,
                          size_t ws_bytes, void *stream);
/* fp32 op-level view of the same computation in the reference's tensor layouts
 * (compute_pixel_in_image_mv + get_projection_features_mv): pix [b,nv,nr,S,2],
 * corner (x0,y0) int32 [b,nv,nr,S,2], gathered [b,nv,nr,S,259] (nullable),
 * cam [b,nv,nr,S,4] camera-frame homogeneous points (nullable).                       */
int cn_project_gather_debug(const CnScene *scene, const float *ray_o, const float *ray_d,
                            const float *z, int32_t nr, int32_t S, float *pix, int32_t *corner,
                            float *gathered, float *cam, void *stream);
/* tensorflow_addons.image.interpolate_bilinear(grid [B,H,W,C], query [B,N,2], 'xy')
 * -> out [B,N,C]  (call site nerf_utils.py:282; edge clamp, alpha clamp)              */
int cn_interpolate_bilinear(const float *grid, const float *query, int32_t B, int32_t H, int32_t W,
                            int32_t C, int64_t N, float *out, void *stream);
/* position_encoding (nerf_utils.py:108-126): pos [n,D] -> out [n, D*n_freq*2]         */
int cn_position_encoding(const float *pos, int64_t n, int32_t D, int32_t n_freq, float freq0,
                         float *out, void *stream);
/* backward: dX tile images -> d_features [b,nv,H,W,256] (nullable, +=) and dz [b,nr,S]
 * (nullable, +=) through bilinear coordinates and posenc(p_cam).  `ws` must hold the records
 * written by cn_project_gather_fwd / cn_project_bin for the same scene, rays and z.            */
int cn_project_gather_bwd(const CnScene *scene, const float *ray_o, const float *ray_d,
                          const float *z, int32_t nr, int32_t S, const void *dx_tiles, void *ws,
                          size_t ws_bytes, float *d_features, float *dz, void *stream);

/* Same backward with the gradient w.r.t. the rays themselves: d_ray_o [b,nr,3] and d_ray_d [b,nr,3]
 * (both +=, caller zero-fills; dz nullable, +=).  x = o + z d enters through the projection / posenc(p_cam)
 * chain, d also through posenc(E^-1 [d;1]) (nerf_utils.py:84-105).  This is what the pose-query models need
 * (DeltaNGF / LanguageNeRF / GraspMVNeRF `_call`, delta_ngf/model.py:196-232, and the pose gradient taken by
 * DNGFOptimizer.optimize_pose, grasp_optimizer.py:158-184): a query point is a ray with S = 1, z = 0,
 * origin = point (`poses[..., :3, 3]`), direction = approach axis (`poses[..., :3, :3] . z_dir`).        */
int cn_project_gather_bwd_rays(const CnScene *scene, const float *ray_o, const float *ray_d,
                               const float *z, int32_t nr, int32_t S, const void *dx_tiles, void *ws,
                               size_t ws_bytes, float *dz, float *d_ray_o, float *d_ray_d, void *stream);

/* Feature-map gradient of several passes in ONE sweep (e.g. coarse + fine): every texel of
 * d_features [b,nv,H,W,256] sums the rows of its four neighbouring cells of every set and is
 * written exactly once (accumulate = 0: the map needs no zero fill; accumulate = 1: +=).
 * Host arrays of length n_sets (<= 4): nr, S, ws, ws_bytes, dx_tiles; every workspace must have
 * been filled by cn_project_gather_fwd / cn_project_bin for this scene.
 * feat_up == 2 (n_sets <= 3): the rows of one low-resolution cell row are summed once into a sliding 3x3
 * register window and added to d_features [b,nv,H/2,W/2,256] with red.global (accumulate = 0 zero-fills first). */
int cn_feature_grad(const CnScene *scene, int32_t n_sets, const int32_t *nr_host, const int32_t *S_host,
                    void *const *ws_host, const size_t *ws_bytes_host, const void *const *dx_tiles_host,
                    float *d_features, int32_t accumulate, void *stream);

/* ---- full-image bookkeeping of render_view (model_v0.py:243-281) ------------------------- */

/* get_rays (nerf_utils.py:15-35) for pixels [first_pixel, first_pixel + n) of a W-wide image, row-major
 * (u = column, v = row): ray_d [n,3] = normalise(rk_inv . [u, v, 1]), ray_o [n,3] = origin.
 *   rk_inv_host [9]: pose[:3,:3] @ inv(K[:3,:3]) computed by the caller on the HOST (row-major, float64 storage);
 *   single_precision: 1 = evaluate in float32 (numpy's arithmetic when pose and K are float32), 0 = float64. */
int cn_generate_rays(const double *rk_inv_host, const float *origin_host, int32_t single_precision, int32_t W,
                     int64_t first_pixel, int64_t n, float *ray_o, float *ray_d, void *stream);
/* uint8 conversion of a rendered image (model_v0.py:272-280): rgb8 [n,3] = uint8(clip(rgb * 255, 0, 255)) by
 * truncation, depth8 [n] = uint8((depth - min) / (max - min) * 255); minmax_ws: 8 bytes of device scratch. */
int cn_quantize_render(const float *rgb, const float *depth, int64_t n, void *minmax_ws, uint8_t *rgb8,
                       uint8_t *depth8, void *stream);

/* source images of MVNeRFDataGenerator.get_input (data_generator/mvnerf.py:37-38: `np.array(colors) / 255.0` cast to
 * float32) from the dataset's uint8 pixels: dst [n_pixels,3] = float32(double(src[p, c]) / 255.0), c < 3 of
 * src_channels (3 = RGB, 4 = RGBA as stored).  Lets the training loop upload uint8 images and keeps the host out
 * of the per-step float conversion; bit-identical to the numpy expression.                                        */
int cn_images_u8_to_f32(const uint8_t *src, int64_t n_pixels, int32_t src_channels, float *dst, void *stream);

/* ---- compositing --------------------------------------------------------------- */

/* volumetric_render (model_v0.py:89-100): z [n,S], rgbs [n,S,4] = (r,g,b,sigma)
 * -> rgb [n,3], depth [n], weights [n,S].  S in {64,128}.                              */
int cn_composite_fwd(const float *z, const float *rgbs, int64_t n_rays, int32_t S, float *rgb,
                     float *depth, float *weights, void *stream);
/* backward; transmittance is recomputed.  d_depth, d_weights, dz nullable.
 * d_y [n,S,4] receives the gradient w.r.t. the readout pre-activations
 * (sigmoid / softplus derivatives folded in, layers.py:395-396).                       */
int cn_composite_bwd(const float *z, const float *rgbs, int64_t n_rays, int32_t S,
                     const float *d_rgb, const float *d_depth, const float *d_weights,
                     float *d_y, float *dz, void *stream);

/* ---- MLP (tcgen05 / TMEM) ------------------------------------------------------- */

/* Flat fp32 parameter layout of one net (Keras variable order, kernel[in][out]):
 *   layer_0.kernel [379,128], layer_0.bias [128],
 *   for blk in feature_blocks[0..2], fusion_blocks[0..2]:
 *       layer_0.kernel [128,128], layer_0.bias, layer_1.kernel [128,128], layer_1.bias
 *   output_layer.kernel [128,4], output_layer.bias [4]          (CN_NET_PARAMS floats)
 * cn_pack_weights writes the bf16 operand images + bias tables (CN_PACKED_BYTES).      */
int cn_pack_weights(const float *params, void *packed, void *stream);

/* per-view stage: Dense 379->128 + 3 ResNet blocks (layers.py:363-365).
 *   x_tiles -> h3_tiles [n_tiles] (bf16 activation tiles).
 *   acts (nullable): saved block operands for the backward,
 *   [n_tiles][3] activation tiles relu(h) (one per block; relu(t) is recomputed).       */
int cn_mlp_feature_fwd(const void *packed, const void *x_tiles, int64_t n_tiles, void *h3_tiles,
                       void *acts, void *stream);
/* fusion stage: mean over views + 3 ResNet blocks + RenderReadout
 * (layers.py:366-376,392-397).  h3_tiles is [b][nv][tiles_per_view]; output
 * rgbs [b*tiles_per_view*128, 4] = (sigmoid rgb, softplus sigma).
 *   acts (nullable): [n_out_tiles][4] activation tiles (3 x relu(h), relu(e)).           */
int cn_mlp_fusion_fwd(const void *packed, const void *h3_tiles, int32_t b, int32_t nv,
                      int64_t tiles_per_view, float *rgbs, void *acts, void *stream);

/* ONE source view (nv = 1: nerf_1_view, the reference's default config; the pose-query validation loop): the mean
 * over views (layers.py:368-370) is the identity, so both stages and the readout run in one launch --
 * cn_mlp_feature_fwd + cn_mlp_fusion_fwd[_taps / _tap_tiles] (+ the _ex options) with bit-identical results; the bf16
 * per-view output never reaches memory.  n_tiles = b * nr * S / 128.  acts_f / acts_u (both or neither): the saved
 * activation tiles of the two stages in the layouts above (3 and 4 tiles per tile; 6 and 7 with save_hidden).
 * taps / tap_mode: 0 none, 1 fp32 rows [4][n_tiles*128][128], 2 bf16 tile images [4][n_tiles][32 KiB].
 * mask_acts_f / mask_acts_u (both or neither, acts must be NULL): the tangent pass.               */
int cn_mlp_chain_fwd(const void *packed, const void *x_tiles, int64_t n_tiles, float *rgbs, void *acts_f, void *acts_u,
                     void *taps, int32_t tap_mode, int32_t save_hidden, const void *mask_acts_f, const void *mask_acts_u,
                     void *stream);

/* fusion stage with `complete_output` taps (layers.py:352,376-377; consumed as
 * `fine_embedding(...)[4:]` by delta_ngf/model.py:240-241, lmvnerf/model_v4.py, grasp_mvnerf/model.py):
 * taps [4][b*tiles_per_view*128][128] fp32 = (mean over views, output of fusion block 1, 2, 3), pre-activation.
 * Everything else as cn_mlp_fusion_fwd.                                                       */
int cn_mlp_fusion_fwd_taps(const void *packed, const void *h3_tiles, int32_t b, int32_t nv,
                           int64_t tiles_per_view, float *rgbs, void *acts, float *taps, void *stream);
/* g_tiles [n_tiles] (bf16 activation-tile images) = (accumulate ? g_tiles : 0) + rows [n_tiles*128][128] fp32:
 * the gradient of a tapped stage output joins the gradient stream of cn_mlp_block_bwd.        */
int cn_tile_add_rows(void *g_tiles, const float *rows, int64_t n_tiles, int32_t accumulate, void *stream);
/* cn_mlp_fusion_fwd_taps with the four stage outputs written as bf16 activation-tile images
 * tap_tiles [4][b*tiles_per_view][32 KiB] -- the operand tiles cn_grasp_head_fwd consumes (half the bytes of the
 * fp32 rows, no conversion pass).                                                               */
int cn_mlp_fusion_fwd_tap_tiles(const void *packed, const void *h3_tiles, int32_t b, int32_t nv,
                                int64_t tiles_per_view, float *rgbs, void *acts, void *tap_tiles, void *stream);
/* g_tiles [n_tiles] += src_tiles [n_tiles] (both bf16 activation-tile images): the tap gradients cn_grasp_head_bwd
 * returns as tiles join the gradient stream without an fp32 round trip.                        */
int cn_tile_add_tiles(void *g_tiles, const void *src_tiles, int64_t n_tiles, void *stream);

/* ---- tangent (forward-mode) pass of the pose-query embedding: the double backward of DeltaNGF.train_step /
 * LanguageNeRF.train_step (delta_ngf/model.py:264-296, lmvnerf/model_v4.py:284-322: tape_1.gradient(prediction,
 * pose_variables) inside the outer tape) differentiates d prediction / d pose w.r.t. the head's weights, which needs
 * J . u of the embedding for a pose-space direction u -- the forward kernels on tangent inputs, with the ReLU patterns
 * of the primal pass (its saved activation tiles) and without biases.
 *   cn_query_tangent_x: points, dirs, u_points, u_dirs [b][n][3] (n a multiple of 128; one sample per "ray") ->
 *     x_tiles [b*nv*n/128][96 KiB], the tangent of the MLP-input rows in the layout of cn_project_gather_fwd.
 *   cn_mlp_feature_fwd_ex / cn_mlp_fusion_fwd_taps_ex: cn_mlp_feature_fwd / cn_mlp_fusion_fwd_taps(_tap_tiles) with
 *     save_hidden = 1: acts holds 3 more tiles per tile, the hidden activations relu(t + ba) of the blocks behind the
 *       usual ones (per-view stage 6 tiles per tile, fusion stage 7: act_stride_tiles of the backward kernels);
 *     mask_acts != NULL (acts must be NULL): the tangent pass -- mask_acts is the acts buffer a save_hidden forward
 *       of the same rows wrote; every ReLU becomes a multiplication by the sign pattern of the saved tile, every
 *       bias is dropped; taps then are the tangents of the four stage outputs (rgbs is scratch).               */
int cn_query_tangent_x(const CnScene *scene, const float *points, const float *dirs, const float *u_points,
                       const float *u_dirs, int32_t n, void *x_tiles, void *stream);
int cn_mlp_feature_fwd_ex(const void *packed, const void *x_tiles, int64_t n_tiles, void *h3_tiles, void *acts,
                          int32_t save_hidden, const void *mask_acts, void *stream);
int cn_mlp_fusion_fwd_taps_ex(const void *packed, const void *h3_tiles, int32_t b, int32_t nv, int64_t tiles_per_view,
                              float *rgbs, void *acts, void *taps, int32_t tap_bf16, int32_t save_hidden,
                              const void *mask_acts, void *stream);

/* poses = transforms @ transforms_to_check (delta_ngf/model.py:199-200, 221-223; same lines in lmvnerf/model_v4.py,
 * grasp_mvnerf/model.py): transforms [n_poses][4][4] (bottom row [0 0 0 1]), transforms_to_check [n5][4][4] ->
 * query points and approach directions [n_poses * n5][3] in '(np n5)' row order, and the backward to the pose
 * matrices (d_transforms [n_poses][4][4], bottom row zero).                                      */
int cn_pose_points_fwd(const float *transforms, const float *transforms_to_check, int64_t n_poses, int32_t n5,
                       float *points, float *dirs, void *stream);
int cn_pose_points_bwd(const float *d_points, const float *d_dirs, const float *transforms_to_check, int64_t n_poses,
                       int32_t n5, float *d_transforms, void *stream);

/* ---- GraspReadout.call (delta_ngf/layers.py:30-42; relu variant grasp_mvnerf/layers.py:22-31) on tcgen05, and its
 * backward to the four stage outputs -- the gradient DNGFOptimizer.optimize_pose takes (grasp_optimizer.py:158-184).
 * Rows are the fusion-stage rows of the pose-query call, '(b (np n5))' padded per scene: row R belongs to scene
 * R / n_pad, i = R % n_pad, pose i / 42 and offset i % 42 of that scene for i < np * 42 (padding rows are computed and
 * dropped).  n_tiles = b * n_pad / 128.  activation: 1 = elu, 0 = relu.
 *   cn_grasp_head_pack: params [CN_HEAD_PARAMS] fp32 -> pack [CN_HEAD_PACK_BYTES] (once per weight version).
 *   cn_grasp_head_fwd : tap_tiles [4][n_tiles][32 KiB] bf16 tile images (cn_mlp_fusion_fwd_tap_tiles, or fp32 rows
 *     through cn_tile_add_rows) -> y [b*np] fp32.  Kept for the
 *     backward (caller-allocated, P = ceil(b*np / 128) pose tiles): vslabs [P][42][2][16 KiB] (the flattened
 *     'b np (n5 d)' activations and their second activation as bf16 K-chunks), acat (nullable) [n_tiles][64 KiB],
 *     saved [P][80 KiB] (activation outputs of the per-pose tail).
 *   cn_grasp_head_bwd : gy [b*np] fp32 -> dtap_tiles [4][n_tiles] bf16 activation-tile images (d y / d tap_s);
 *     dpc_ws [P][42][16 KiB] is scratch.  No parameter gradients: training the head goes through the differentiable
 *     torch form (src/lib/delta_ngf/layers.py), which the double backward of DeltaNGF.train_step needs anyway.   */
int cn_grasp_head_pack(const float *params, void *pack, void *stream);
int cn_grasp_head_fwd(const void *tap_tiles, int64_t n_tiles, int32_t b, int32_t n_pad, int32_t np, const void *pack,
                      int32_t activation, void *vslabs, void *acat, void *saved, float *y, void *stream);
int cn_grasp_head_bwd(const float *gy, int64_t n_tiles, int32_t b, int32_t n_pad, int32_t np, const void *pack,
                      int32_t activation, const void *vslabs, const void *acat, const void *saved, void *dpc_ws,
                      void *dtap_tiles, void *stream);

/* cn_grasp_head_bwd without the dpc_ws argument.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  The scratch is a file-scope `static __device__` array inside the library, sized
 * for at most CN_HEAD_BWD_STATIC_MAX_POSES poses (b * np; 8 pose tiles x 42 x 16 KiB = 5.25 MiB, held by every GPU that
 * loads the library until the process exits); more poses are rejected with an error (cn_last_error) and nothing is
 * launched.  The call allocates nothing (no cudaMalloc / cudaFree); inputs and outputs are the caller's, as above, and
 * the results are those of cn_grasp_head_bwd bit for bit (same kernels).  ALL calls on a device share the one array:
 * calls that may overlap on the GPU (different streams without an ordering between them) race on it -- order them,
 * or use cn_grasp_head_bwd with one dpc_ws per stream.  The array keeps the previous call's contents; nothing in it
 * needs clearing (every row that is read is written first by the same call).  The first call on each device resolves
 * the array's device address (cudaGetSymbolAddress, with that device current): make it at set-up, outside capture. */
#define CN_HEAD_BWD_STATIC_MAX_POSES 1024
int cn_grasp_head_bwd_static(const float *gy, int64_t n_tiles, int32_t b, int32_t n_pad, int32_t np, const void *pack,
                             int32_t activation, const void *vslabs, const void *acat, const void *saved,
                             void *dtap_tiles, void *stream);

/* readout backward: d_y [rows,4] + relu(e) tiles -> g tiles (bf16), d(output_layer)
 * accumulated into grad (flat layout above).                                           */
int cn_mlp_readout_bwd(const float *params, const float *d_y, const void *acts, int32_t act_stride_tiles,
                       int32_t act_index, int64_t n_tiles, void *g_tiles, float *grad, void *stream);
/* backward of one ResNet block for every tile: g_y -> g_h (in place allowed),
 * dW/db of both Dense layers accumulated into grad.  blk in 0..5.  acts tile
 * [act_index] of every tile is the block's saved relu(h); the hidden activation is
 * recomputed on the tensor cores (bit-identical to the forward).
 * The mean over views (layers.py:368-370) is folded in: with gy_views = nv > 1, tile
 * (b, v, t) reads the fusion-stage gradient tile (b, t) of gy_tiles [b][gy_tiles_per_view]
 * (gh_tiles must then be a separate [n_tiles] buffer) and `scale` = 1/nv multiplies what is
 * accumulated into grad, in fp32 (the bf16 g stream itself stays unscaled; every later
 * per-view stage kernel takes the same scale).  gy_views = 1, scale = 1: plain block.
 * grad == NULL: only g_h is formed (no weight-gradient GEMMs) -- the backward of a frozen
 * field to the query poses (grasp_optimizer.py:158-184).                                 */
int cn_mlp_block_bwd(const void *packed, int32_t blk, const void *acts, int32_t act_stride_tiles,
                     int32_t act_index, int64_t n_tiles, const void *gy_tiles, int32_t gy_views,
                     int64_t gy_tiles_per_view, void *gh_tiles, float scale, float *grad, void *stream);
/* Backward of a whole MLP stage in ONE launch (layers.py:363-375 in reverse): stage 1 = fusion blocks 5, 4, 3,
 * stage 0 = per-view blocks 2, 1, 0 and -- when x_tiles / dx_tiles are given -- the first Dense's weight gradient
 * (cn_mlp_l0_wgrad) and input gradient (cn_mlp_l0_dgrad).  Role-specialised persistent CTAs (one role per block, one
 * each for the first Dense's dgrad / wgrad) pass the gradient stream from role to role through small tile rings in
 * `ws` that stay in L2, so g never makes the HBM round trips it makes between the per-block launches; results equal
 * the chain of the per-block calls (the bf16 g stream bit for bit, the fp32 sums up to summation order).
 *   acts             [n_tiles][act_stride_tiles] saved relu(h) tiles, tile index 0..2 = the stage's blocks in order
 *   gy_tiles         gradient w.r.t. the stage output; gy_views / gy_tiles_per_view as in cn_mlp_block_bwd
 *   gh_tiles         gradient w.r.t. the stage input (may alias gy_tiles when gy_views == 1), or NULL when the
 *                    first Dense roles consume it (stage 0 with x_tiles and / or dx_tiles; it is then not stored)
 *   x_tiles, dx_tiles  nullable: MLP-input tiles (-> dW0, db0 accumulated into grad) / their gradient (written)
 *   ws               cn_mlp_stage_bwd_ws_bytes() bytes of device scratch, 256-byte aligned (rings + counters)     */
size_t cn_mlp_stage_bwd_ws_bytes(void);
int cn_mlp_stage_bwd(const void *packed, int32_t stage, const void *acts, int32_t act_stride_tiles, int64_t n_tiles,
                     const void *gy_tiles, int32_t gy_views, int64_t gy_tiles_per_view, void *gh_tiles,
                     const void *x_tiles, void *dx_tiles, float scale, float *grad, void *ws, size_t ws_bytes,
                     void *stream);
/* mean-over-views backward: g_m [b][tiles_per_view] -> g_h3 [b][nv][tiles_per_view] = g_m / nv */
int cn_mlp_mean_bwd(const void *gm_tiles, int32_t b, int32_t nv, int64_t tiles_per_view,
                    void *gh3_tiles, void *stream);
/* first Dense backward: dX = scale * g0 . W0^T (tile images, same column order as X) ... */
int cn_mlp_l0_dgrad(const void *packed, const void *g0_tiles, int64_t n_tiles, void *dx_tiles,
                    float scale, void *stream);
/* ... and dW0 += scale * X^T . g0, db0 += scale * sum(g0), accumulated into grad.      */
int cn_mlp_l0_wgrad(const void *x_tiles, const void *g0_tiles, int64_t n_tiles, float scale,
                    float *grad, void *stream);

/* ---- parameter update ----------------------------------------------------------- */

/* The update MVVNeRFRenderer.train_step applies (model_v0.py:186-197) as one launch over flat fp32 buffers:
 * optimize() (nerf_utils.py:8-12: per-element clip_by_value(g, -clip, clip), clip <= 0 = off) followed by
 * tf.keras.optimizers.Adam(learning_rate=WarmupScheduler(target_lr, warmup_steps, scale_down_after))
 * (train_nerf.py:22-23, nerf_utils.py:288-300):
 *   t = *step_counter + 1;  lr = t <= warmup ? t / warmup * target : (t <= scale_down_after ? target : 0.1 target)
 *   g = clip(grad * grad_scale);  m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2
 *   p -= lr sqrt(1 - b2^t) / (1 - b1^t) * m / (sqrt(v) + eps);   *step_counter = t      (grad is left clipped)
 * The buffers are [n_rows][row_stride]; the first n_update elements of every row are updated (the reference's
 * optimizer groups hold the two embeddings only, train_nerf.py:27-32: n_update = CN_EMB_PARAMS leaves the readouts
 * untouched).  step_counter is a DEVICE double, so a CUDA graph that recorded the launch replays the schedule.      */
int cn_adam_step(float *params, float *grad, float *m, float *v, int64_t n_rows, int64_t row_stride,
                 int64_t n_update, double *step_counter, double target_lr, double warmup_steps,
                 double scale_down_after, float grad_scale, float clip, double beta1, double beta2, double eps,
                 void *stream);

/* ---- diagnostics ---------------------------------------------------------------- */

/* One 128 x 128 x K tcgen05.mma tile with caller-supplied operand images and
 * descriptor fields; used by tests to pin the shared-memory descriptor encoding.       */
int cn_probe_umma(const void *a_img, int32_t a_bytes, const void *b_img, int32_t b_bytes,
                  int32_t n_mma, int32_t a_major_mn, int32_t b_major_mn, int32_t a_lbo,
                  int32_t a_sbo, int32_t a_kstep, int32_t b_lbo, int32_t b_sbo, int32_t b_kstep,
                  float *d_out, void *stream);

/* Throughput probe for the two ways of adding a 9x9-texel tile (81 KiB) to a [n_img,H,W,256]
 * fp32 map: mode 0 = cp.reduce.async.bulk.add.f32 from shared memory, mode 1 =
 * red.global.add.v4.f32 from registers.  Every CTA adds `iters` tiles at pseudo-random
 * positions; the caller times the launch.                                                      */
int cn_probe_tile_reduce(float *map, int32_t n_img, int32_t H, int32_t W, int32_t iters, int32_t mode,
                         int32_t ctas_per_sm, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* CLIPNERF_B200_H */
