"""CPU oracle for the CLIP-NeRF volumetric render hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import this
module; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` do, and only as the checker / the timed
CPU baseline.

What this is
------------
An op-for-op PyTorch (CPU, fp32 by default, fp64 on request) restatement of the
TensorFlow/Keras reference path ``MVVNeRFRenderer._call`` and of every function it
calls.  Each function cites the reference file:line it follows
(paths relative to the reference checkout root).

Parity status: the reference cannot run here (TensorFlow, tensorflow_addons are
not installed; no network) and it ships no tests / golden vectors.  The
restatement is pinned in two ways:
  * ``oracle/tf_shim`` executes the *reference's own source files* on top of a
    small torch-backed ``tensorflow`` stand-in and ``tests/test_oracle_vs_reference_source.py``
    compares this module against them (run in the build container, where
    ``/root/reference`` exists); golden vectors produced that way are committed
    under ``tests/golden/``.
  * closed-form unit tests and fp64 ``gradcheck``.
Third-party arithmetic that is *not* in the reference tree
(``tensorflow_addons.image.interpolate_bilinear``, TF kernels' reduction order)
is restated from its published algorithm; where TF leaves the order of a
floating-point reduction unspecified this oracle pins a sequential order and
says so at the call site.

Random numbers are explicit inputs (``u_strat``, ``u_pdf``) so that the CUDA path
can be compared bit-for-bit on indices.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor

# --------------------------------------------------------------------------
# Host-side ray / camera bookkeeping (numpy in the reference as well)
# --------------------------------------------------------------------------


def get_specific_rays(u, v, extrinsics, intrinsics, norm_direction_vector=True):
    """``src/lib/mvnerf/nerf_utils.py:27-35``: d = R . K^-1 . [u, v, 1], o = pose[:3, 3]."""
    pixels = np.stack((u, v, np.ones_like(u)), axis=0)
    rays_d = extrinsics[:3, :3] @ np.linalg.inv(intrinsics[:3, :3]) @ pixels
    rays_d = rays_d.T
    if norm_direction_vector:
        rays_d = rays_d / np.linalg.norm(rays_d, axis=1, keepdims=True)
    rays_o = np.broadcast_to(extrinsics[:3, -1], rays_d.shape)
    return rays_o, rays_d


def get_rays(image_width, image_height, extrinsics, intrinsics, norm_direction_vector=True):
    """``nerf_utils.py:15-24``: row-major pixel grid, u = column, v = row."""
    u, v = np.meshgrid(np.arange(image_width, dtype=np.float32),
                       np.arange(image_height, dtype=np.float32), indexing='xy')
    u = u.reshape(-1)
    v = v.reshape(-1)
    rays_o, rays_d = get_specific_rays(u, v, extrinsics, intrinsics, norm_direction_vector)
    rays_d = rays_d.reshape(image_height, image_width, 3)
    rays_o = rays_o.reshape(image_height, image_width, 3)
    return rays_o, rays_d


def bbox_biased_sample(n_sample, bboxes, image_height, image_width, in_box_p=0.8, rng=None):
    """``nerf_utils.py:38-46``.  The reference draws from numpy's *global* RNG
    (``np.random.randint``); pass ``rng=None`` to do the same, or a
    ``np.random.RandomState`` for a private stream with identical draws."""
    r = np.random if rng is None else rng
    n_inside = int(n_sample * in_box_p)
    n_random = n_sample - n_inside
    in_samples = r.randint(bboxes[:2], bboxes[2:], (n_inside, 2))
    random_samples = r.randint((image_height, image_width), size=(n_random, 2))
    return np.concatenate([in_samples, random_samples], axis=0)


def camera_parameters(camera_config):
    """``src/lib/data_generator/util.py:4-10``: K 3x3 -> 4x4, E^-1 = inv(pose)."""
    src_extrinsic = camera_config['pose']
    src_intrinsic = np.reshape(camera_config['intrinsics'], (3, 3))
    src_intrinsic = np.concatenate((src_intrinsic, np.zeros((3, 1))), axis=1)
    src_intrinsic = np.concatenate((src_intrinsic, np.array([[0, 0, 0, 1]])), axis=0)
    src_extrinsic_inv = np.linalg.inv(src_extrinsic)
    return src_extrinsic_inv, src_intrinsic


def get_input(colors, camera_configs, r_d, r_o):
    """``src/lib/data_generator/mvnerf.py:28-43`` (5-tuple of fp32 arrays, batch of 1)."""
    e_invs, ks = [], []
    for cc in camera_configs:
        e, k = camera_parameters(cc)
        e_invs.append(e)
        ks.append(k)
    return (np.array([r_o], dtype=np.float32),
            np.array([r_d], dtype=np.float32),
            np.array([np.array(colors) / 255.0], dtype=np.float32),
            np.array([ks], dtype=np.float32),
            np.array([e_invs], dtype=np.float32))


def get_target(color, rays):
    """``mvnerf.py:46-48``: rays are (row, col) ints."""
    return np.array(color[rays[:, 0], rays[:, 1], :3]) / 255.0


def quantize_render(all_rgbs: np.ndarray, all_depths: np.ndarray, image_shape):
    """``src/lib/mvnerf/model_v0.py:275-280``: x255, clip, *truncating* uint8 cast;
    depth min-max normalised then truncated."""
    rgbs = np.reshape(np.array(all_rgbs), (*image_shape, 3)) * 255
    rgbs = np.clip(rgbs, 0, 255).astype(np.uint8)
    depths = np.reshape(np.array(all_depths), (*image_shape, 1))
    norm = (depths - np.min(depths)) / (np.max(depths) - np.min(depths))
    depths = (norm * 255).astype(np.uint8)
    return rgbs, depths


# --------------------------------------------------------------------------
# Tensor ops of the hot path
# --------------------------------------------------------------------------


def sample_along_ray(rays_origin: Tensor, rays_direction: Tensor, near: float, far: float,
                     n_samples: int, u_strat: Tensor) -> Tuple[Tensor, Tensor]:
    """``nerf_utils.py:49-61``.  ``u_strat`` [b, nr, S] replaces ``tf.random.uniform``.

    The interval table is built in Python doubles and cast once (``:51``);
    ``z = 1 * lo + u * step`` with separately rounded product and sum (no FMA).
    """
    step = (far - near) / n_samples
    dt = rays_origin.dtype
    lower = torch.tensor([near + i * step for i in range(n_samples + 1)], dtype=torch.float32).to(dt)[:-1]
    lower_bound = torch.ones_like(rays_direction[..., 0])[..., None] * lower
    random_added = u_strat * step
    z = lower_bound + random_added
    world_points = rays_origin[:, :, None, :] + z[..., None] * rays_direction[:, :, None, :]
    return world_points, z


def _affine_rows(M: Tensor, x: Tensor, y: Tensor, z: Tensor, w, rows: int) -> List[Tensor]:
    """rows of M . [x, y, z, w] as explicitly ordered fp32 products and sums
    ((m0 x + m1 y) + m2 z) + m3 w, each operation rounded on its own.

    ``tf.matmul`` (cuBLAS batched 4x4 GEMMs, ``nerf_utils.py:71,74,100``) leaves the order of the
    four-term dot product unspecified; the oracle pins this one so that pixel coordinates and
    bilinear corner indices can be compared bit-for-bit."""
    out = []
    for i in range(rows):
        acc = M[..., i, 0] * x
        acc = acc + M[..., i, 1] * y
        acc = acc + M[..., i, 2] * z
        acc = acc + M[..., i, 3] * w
        out.append(acc)
    return out


def compute_pixel_in_image_mv(world_points: Tensor, src_intrinsics: Tensor,
                              src_extrinsics_inv: Tensor) -> Tuple[Tensor, Tensor]:
    """``nerf_utils.py:64-81``: p = E^-1 [x,1]; q = K p; pix = clip(q.xy / max(q.z, 1e-8), +-1e6).

    Returns pix [b, nv, nr, S, 2] and camera points [b, nv, nr, S, 4].
    """
    x = world_points[:, None, :, :, 0]
    y = world_points[:, None, :, :, 1]
    z = world_points[:, None, :, :, 2]
    E = src_extrinsics_inv[:, :, None, None]                                # [b, nv, 1, 1, 4, 4]
    K = src_intrinsics[:, :, None, None]
    p = _affine_rows(E, x, y, z, 1.0, 4)
    cam = torch.stack(p, dim=-1)                                            # [b, nv, nr, S, 4]
    q = _affine_rows(K, p[0], p[1], p[2], p[3], 3)
    den = torch.clamp(q[2], min=1e-8)
    pix = torch.stack([q[0] / den, q[1] / den], dim=-1)
    pix = torch.clamp(pix, min=-1e6, max=1e6)
    return pix, cam


def world_to_camera_direction_vector_mv(world_dirs: Tensor, extrinsics_inverse: Tensor) -> Tensor:
    """``nerf_utils.py:84-105``: homogeneous **w = 1** (quirk Q1) so the result is R d + t."""
    E = extrinsics_inverse[:, :, None]                                      # [b, nv, 1, 4, 4]
    d = world_dirs[:, None]                                                 # [b, 1, nr, 3]
    rows = _affine_rows(E, d[..., 0], d[..., 1], d[..., 2], 1.0, 3)
    return torch.stack(rows, dim=-1)                                        # [b, nv, nr, 3]


def interpolate_bilinear(grid: Tensor, query_points: Tensor, indexing: str = 'xy') -> Tensor:
    """Restatement of ``tensorflow_addons.image.interpolate_bilinear`` (TFA <= 0.20,
    ``tensorflow_addons/image/dense_image_warp.py``; called at ``nerf_utils.py:282``).

    grid [B, H, W, C], query_points [B, N, 2] -> [B, N, C].  Per axis:
    ``floor = clamp(floor(q), 0, size-2)``, ``alpha = clamp(q - floor, 0, 1)``;
    ``top = ax*(TR-TL)+TL``, ``bottom = ax*(BR-BL)+BL``, ``out = ay*(bottom-top)+top``.
    """
    B, H, W, C = grid.shape
    if indexing == 'xy':
        qx, qy = query_points[..., 0], query_points[..., 1]
    else:
        qy, qx = query_points[..., 0], query_points[..., 1]

    def axis(q, size):
        max_floor = float(size - 2)
        fl = torch.clamp(torch.floor(q), min=0.0, max=max_floor)   # tf.minimum(tf.maximum(0, floor(q)), size-2)
        int_fl = fl.detach().to(torch.int64)
        alpha = torch.clamp(q - fl.detach(), min=0.0, max=1.0)
        return int_fl, alpha

    fy, ay = axis(qy, H)
    fx, ax = axis(qx, W)
    flat = grid.reshape(B, H * W, C)

    def gather(yc, xc):
        lin = (yc * W + xc)[..., None].expand(-1, -1, C)
        return torch.gather(flat, 1, lin)

    tl = gather(fy, fx)
    tr = gather(fy, fx + 1)
    bl = gather(fy + 1, fx)
    br = gather(fy + 1, fx + 1)
    ax = ax[..., None]
    ay = ay[..., None]
    top = ax * (tr - tl) + tl
    bottom = ax * (br - bl) + bl
    return ay * (bottom - top) + top


def upsample2x_bilinear(x: Tensor) -> Tensor:
    """``tf.keras.layers.UpSampling2D(size=(2, 2), interpolation='bilinear')`` -- the last layer of every
    CombineCLIPVisual variant (``layers.py:547,589``, ``617,659``; ``legacy_layers.py:190``; ``model_without.py:28,76``).

    Keras forwards to ``tf.image.resize(..., method='bilinear')``; TensorFlow (third-party, absent here; the
    reference pins tensorflow 2.x in its environment) computes it with half-pixel centres and no antialiasing
    (``ResizeBilinear`` kernel, ``half_pixel_centers=True``).  Per axis, for output index o of an input of
    ``size`` elements:

        in    = (o + 0.5) * 0.5 - 0.5
        lower = max(floor(in), 0),  upper = min(ceil(in), size - 1),  lerp = in - floor(in)

    and per pixel ``top = TL + (TR - TL) * x_lerp``, ``bottom = BL + (BR - BL) * x_lerp``,
    ``out = top + (bottom - top) * y_lerp``.  x [..., H, W, C] -> [..., 2H, 2W, C].
    ``tests/test_oracle.py`` pins this against ``torch.nn.functional.interpolate(scale_factor=2, mode='bilinear',
    align_corners=False)``, an independent implementation of the same convention.
    """
    H, W = x.shape[-3], x.shape[-2]

    def axis(size):
        o = torch.arange(2 * size, dtype=torch.float64)
        src = (o + 0.5) * 0.5 - 0.5
        fl = torch.floor(src)
        lower = torch.clamp(fl, min=0).to(torch.int64)
        upper = torch.clamp(torch.ceil(src), max=size - 1).to(torch.int64)
        return lower, upper, (src - fl).to(x.dtype)                 # lerp is 0.25 or 0.75, exact in any float type

    ly, uy, wy = axis(H)
    lx, ux, wx = axis(W)
    wx = wx[:, None]
    wy = wy[:, None, None]
    rows_l, rows_u = x.index_select(-3, ly), x.index_select(-3, uy)

    def x_pass(r):
        left, right = r.index_select(-2, lx), r.index_select(-2, ux)
        return left + (right - left) * wx

    top, bottom = x_pass(rows_l), x_pass(rows_u)
    return top + (bottom - top) * wy


def lowres_stencil(f0: Tensor, alpha: Tensor, size_lr: int):
    """One axis of ``interpolate_bilinear(upsample2x_bilinear(x))`` written as taps on the LOW-resolution axis.

    ``f0`` (int64) / ``alpha`` are the clamped floor and weight of ``interpolate_bilinear`` on the 2 * size_lr axis
    (f0 in [0, 2 size_lr - 2]).  With k = f0 >> 1 the two upsampled texels U[f0], U[f0+1] are

        f0 even:  U[2k]   = 0.25 x[k-1] + 0.75 x[k],   U[2k+1] = 0.75 x[k] + 0.25 x[k+1]
        f0 odd :  U[2k+1] = 0.75 x[k] + 0.25 x[k+1],   U[2k+2] = 0.25 x[k] + 0.75 x[k+1]

    (indices clamped to [0, size_lr - 1]: the resize clamps ``lower`` / ``upper`` the same way), so
    (1 - a) U[f0] + a U[f0+1] = sum_j w_j x[clamp(k + j - 1)] with

        even: w = (0.25 (1 - a), 0.75, 0.25 a)        odd: w = (0, 0.75 - 0.5 a, 0.25 + 0.5 a)

    Returns (idx [..., 3] int64, w [..., 3], dw/da [..., 3]).
    """
    k = f0 >> 1
    odd = (f0 & 1).to(alpha.dtype)
    even = 1.0 - odd
    w = torch.stack([even * 0.25 * (1.0 - alpha), even * 0.75 + odd * (0.75 - 0.5 * alpha),
                     even * 0.25 * alpha + odd * (0.25 + 0.5 * alpha)], dim=-1)
    dw = torch.stack([even * -0.25, odd * -0.5, even * 0.25 + odd * 0.5], dim=-1).expand_as(w)
    idx = torch.stack([k - 1, k, k + 1], dim=-1).clamp(0, size_lr - 1)
    return idx, w, dw


def interpolate_bilinear_lowres(grid_lr: Tensor, query_points: Tensor, indexing: str = 'xy') -> Tensor:
    """``interpolate_bilinear(upsample2x_bilinear(grid_lr), query_points)`` without the upsampled map: the 2x2
    footprint on the [2H, 2W] map is the 3x3 (at most) stencil of ``lowres_stencil`` on grid_lr [B, H, W, C].
    This is the form the CUDA path evaluates when ``CnScene.feat_up == 2``; ``tests/test_oracle.py`` proves it equal
    to the two-stage form (fp64: 1e-13, fp32: 1e-6 of the map's range)."""
    B, H, W, C = grid_lr.shape
    if indexing == 'xy':
        qx, qy = query_points[..., 0], query_points[..., 1]
    else:
        qy, qx = query_points[..., 0], query_points[..., 1]

    def axis(q, size_hr):
        fl = torch.clamp(torch.floor(q), min=0.0, max=float(size_hr - 2))
        return fl.detach().to(torch.int64), torch.clamp(q - fl.detach(), min=0.0, max=1.0)

    fy, ay = axis(qy, 2 * H)
    fx, ax = axis(qx, 2 * W)
    iy, wy, _ = lowres_stencil(fy, ay, H)
    ix, wx, _ = lowres_stencil(fx, ax, W)
    flat = grid_lr.reshape(B, H * W, C)
    out = 0
    for i in range(3):
        row = 0
        for j in range(3):
            lin = (iy[..., i] * W + ix[..., j])[..., None].expand(-1, -1, C)
            row = row + wx[..., j, None] * torch.gather(flat, 1, lin)
        out = out + wy[..., i, None] * row
    return out


def get_projection_features_mv(inputs: Tensor, features: Tensor, pixel_locations: Tensor) -> Tensor:
    """``nerf_utils.py:277-285``: concat(image*2-1, features) (259 ch) and bilinear gather.

    inputs [b, nv, H, W, 3] (already ``2*I-1``), features [b, nv, H, W, C],
    pixel_locations [b, nv, nr, S, 2] -> [b, nv, nr, S, 3+C].
    """
    b, nv, H, W, _ = inputs.shape
    nr, S = pixel_locations.shape[2], pixel_locations.shape[3]
    comb = torch.cat([inputs, features], dim=-1).reshape(b * nv, H, W, -1)
    flat_pix = pixel_locations.reshape(b * nv, nr * S, 2)
    out = interpolate_bilinear(comb, flat_pix, indexing='xy')
    return out.reshape(b, nv, nr, S, -1)


def position_encoding(position: Tensor, n_freq: int = 10, pos_encoding_freq: float = math.pi) -> Tensor:
    """``nerf_utils.py:108-126``: f_k = fp32(pi) * 2^k; per coordinate
    ``[sin f0, cos f0, sin f1, cos f1, ...]``, coordinate-major ('d n f')."""
    dt = position.dtype
    freq = (torch.tensor(pos_encoding_freq, dtype=torch.float32) *
            torch.pow(torch.tensor(2.0, dtype=torch.float32), torch.arange(n_freq, dtype=torch.float32))).to(dt)
    arg = position[..., None] * freq                                       # [..., d, n]
    enc = torch.stack([torch.sin(arg), torch.cos(arg)], dim=-1)            # [..., d, n, 2]
    return enc.reshape(*position.shape[:-1], -1)


def sigma_to_alpha(sigma: Tensor, dists: Tensor) -> Tensor:
    """``nerf_utils.py:129-140``."""
    return 1.0 - torch.exp(-dists * F.relu(sigma))


def volumetric_render(zs: Tensor, density: Tensor, chromacity: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """``model_v0.py:89-100``: last interval repeats the previous one,
    T = exclusive cumprod(1 - alpha + 1e-10), no background term."""
    dists = zs[..., 1:] - zs[..., :-1]
    dists = torch.cat([dists, dists[..., -1:]], dim=-1)
    alpha = sigma_to_alpha(density, dists)
    t_incl = torch.cumprod(1 - alpha + 1e-10, dim=-1)
    transmittance = torch.cat([torch.ones_like(t_incl[..., :1]), t_incl[..., :-1]], dim=-1)
    weights = alpha * transmittance
    rgb = torch.sum(weights[..., None] * chromacity, dim=-2)
    depth = torch.sum(weights * zs, dim=-1)
    return rgb, depth, weights


def _seq_cumsum(x: Tensor) -> Tensor:
    """Inclusive cumulative sum along the last axis as a strict left-to-right chain of additions in
    the tensor's own dtype (``torch.cumsum`` on CPU accumulates fp32 inputs in double)."""
    cols = [x[..., 0]]
    for k in range(1, x.shape[-1]):
        cols.append(cols[-1] + x[..., k])
    return torch.stack(cols, dim=-1)


def sample_pdf(bins: Tensor, weights: Tensor, n_samples: int, u: Tensor,
               return_indices: bool = False):
    """``nerf_utils.py:143-176`` (quirk Q5).  ``u`` [b, nr, n_samples] replaces
    ``tf.random.uniform``.  bins [b, nr, 63], weights [b, nr, 62].

    * ``above = #{k : u >= cdf[k]}`` over the 63-entry cdf -- NOT clamped;
    * an index of 63 reads 0 (TF-GPU ``gather_nd`` out-of-bounds semantics; the
      reference only runs on GPU, ``src/train_nerf.py:86-87``);
    * reduction order: TF does not specify the order of ``reduce_sum`` / ``cumsum``
      on GPU; the oracle pins a **sequential left-to-right fp32 sum** for both
      (``_seq_cumsum``), which is what makes the bin indices reproducible bit-for-bit.
    """
    stable = weights + 1e-5
    w_sum = _seq_cumsum(stable)[..., -1:]
    w_sum = torch.where(torch.abs(w_sum) == 0, torch.ones_like(w_sum), w_sum)
    pdf = stable / w_sum
    cdf = _seq_cumsum(pdf)
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], dim=-1)          # 63 entries
    nb = bins.shape[-1]
    above = (u[..., :, None] >= cdf[..., None, :]).sum(dim=-1)              # int64 in [1, 63]
    below = torch.clamp(above - 1, 0, nb - 1)

    def gather_oob0(src, idx):
        pad = torch.cat([src, torch.zeros_like(src[..., :1])], dim=-1)      # index nb -> 0
        return torch.gather(pad, -1, torch.clamp(idx, 0, nb))

    cdf_a = gather_oob0(cdf, above)
    cdf_b = gather_oob0(cdf, below)
    bins_a = gather_oob0(bins, above)
    bins_b = gather_oob0(bins, below)
    denom = cdf_a - cdf_b
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    t = (u - cdf_b) / denom
    samples = bins_b + t * (bins_a - bins_b)
    if return_indices:
        return samples, above.to(torch.int32), below.to(torch.int32)
    return samples


# --------------------------------------------------------------------------
# MLP (Keras Dense: y = x @ kernel[in, out] + bias)
# --------------------------------------------------------------------------

N_FREQ = 10
HIDDEN = 128
IN_DIM = 3 * 2 * N_FREQ * 2 + 3 + 256          # 60 + 60 + 3 + 256 = 379


def embedding_param_shapes(in_dim: int = IN_DIM, hidden: int = HIDDEN, n_blocks: int = 6):
    """Variable names/shapes of ``MVResNetMLPNeRFEmbedding(6, 128)`` (``layers.py:334-352``)."""
    shapes = [('layer_0.kernel', (in_dim, hidden)), ('layer_0.bias', (hidden,))]
    nf = n_blocks // 2
    for grp, n in (('feature_blocks', nf), ('fusion_blocks', n_blocks - nf)):
        for i in range(n):
            for l in ('layer_0', 'layer_1'):
                shapes.append((f'{grp}.{i}.{l}.kernel', (hidden, hidden)))
                shapes.append((f'{grp}.{i}.{l}.bias', (hidden,)))
    return shapes


def readout_param_shapes(hidden: int = HIDDEN, out: int = 4):
    """``RenderReadout(4)`` (``layers.py:382-397``)."""
    return [('output_layer.kernel', (hidden, out)), ('output_layer.bias', (out,))]


def init_params(seed: int, bias_std: float = 0.0, dtype=torch.float32) -> Dict[str, Tensor]:
    """glorot_uniform kernels / zero biases (Keras defaults); ``bias_std`` > 0 draws
    N(0, bias_std) biases so that parity tests exercise the bias path."""
    g = torch.Generator().manual_seed(seed)
    p: Dict[str, Tensor] = {}
    for net in ('coarse', 'fine'):
        for name, shape in embedding_param_shapes():
            full = f'{net}_embedding.{name}'
            if name.endswith('kernel'):
                lim = math.sqrt(6.0 / (shape[0] + shape[1]))
                p[full] = ((torch.rand(shape, generator=g) * 2 - 1) * lim).to(dtype)
            else:
                p[full] = (torch.randn(shape, generator=g) * bias_std).to(dtype)
        for name, shape in readout_param_shapes():
            full = f'{net}_readout.{name}'
            if name.endswith('kernel'):
                lim = math.sqrt(6.0 / (shape[0] + shape[1]))
                p[full] = ((torch.rand(shape, generator=g) * 2 - 1) * lim).to(dtype)
            else:
                p[full] = (torch.randn(shape, generator=g) * bias_std).to(dtype)
    return p


class _RoundBF16(torch.autograd.Function):
    """round-to-nearest-even to bfloat16 with a straight-through gradient."""

    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).to(x.dtype)

    @staticmethod
    def backward(ctx, g):
        return g


class _RoundTF32(torch.autograd.Function):
    """round to TensorFloat-32 (10 explicit mantissa bits, round-to-nearest-even), straight-through gradient: the
    operand rounding of TensorFlow's default matmul path on Ampere-and-later GPUs (``tf.config.experimental.
    enable_tensor_float_32_execution`` defaults to True), i.e. the arithmetic class the reference itself trains in"""

    @staticmethod
    def forward(ctx, x):
        if x.dtype != torch.float32:
            return x
        i = x.contiguous().view(torch.int32)
        r = (i + 0x0FFF + ((i >> 13) & 1)) & ~0x1FFF
        return r.view(torch.float32)

    @staticmethod
    def backward(ctx, g):
        return g


def _ident(x):
    return x


def _quantizer(quant: Optional[str]):
    """``quant='bf16'`` makes the oracle round exactly where the CUDA path stores bf16 tensor-core
    operands (MLP input rows, weights, ReLU outputs, the per-view stage output); everything else
    (biases, residual stream, accumulation, readout) stays in the working dtype.  ``None`` is the
    reference's plain fp32 arithmetic.  The quantised variant exists to separate implementation
    errors from the (much larger, ReLU-mask-flipping) effect of reduced operand precision."""
    if quant is None:
        return _ident
    if quant == 'bf16':
        return _RoundBF16.apply
    if quant == 'tf32':
        # matmul operands only, as TF32 execution does (the per-view stage output is not a stored bf16 tile there)
        return _RoundTF32.apply
    raise ValueError(f'unknown quant {quant!r}')


def resnet_mlp_block(x: Tensor, p: Dict[str, Tensor], prefix: str, activation: str = 'relu',
                     shortcut_kernel: Optional[Tensor] = None, quant: Optional[str] = None) -> Tensor:
    """``layers.py:262-298``: y = shortcut(x) + act(act(x) W0 + b0) W1 + b1."""
    q = _quantizer(quant)
    if activation == 'relu':
        act = F.relu
    elif activation == 'elu':
        act = F.elu
    else:
        raise ValueError(f'activation {activation} not supported')
    r = q(act(x))
    r = r @ q(p[f'{prefix}.layer_0.kernel']) + p[f'{prefix}.layer_0.bias']
    r = q(act(r))
    r = r @ q(p[f'{prefix}.layer_1.kernel']) + p[f'{prefix}.layer_1.bias']
    sc = x if shortcut_kernel is None else x @ shortcut_kernel
    return sc + r


def mv_embedding_from_x(x: Tensor, p: Dict[str, Tensor], prefix: str, n_views: int,
                        complete_output: bool = False, quant: Optional[str] = None):
    """Everything of ``MVResNetMLPNeRFEmbedding.call`` after the concat (``layers.py:363-379``):
    x [(b nv), ..., 379] -> Dense 379->128, 3 per-view blocks, mean over views, 3 fusion blocks."""
    q = _quantizer(quant)
    x = q(x) @ q(p[f'{prefix}.layer_0.kernel']) + p[f'{prefix}.layer_0.bias']
    outputs = [x]
    for i in range(3):
        outputs.append(resnet_mlp_block(outputs[-1], p, f'{prefix}.feature_blocks.{i}', quant=quant))
    q_store = q if quant == 'bf16' else _ident          # the CUDA path stores the per-view stage output as bf16 tiles
    pre = q_store(outputs[-1]).reshape(-1, n_views, *outputs[-1].shape[1:])
    outputs.append(pre.mean(dim=1))
    for i in range(3):
        outputs.append(resnet_mlp_block(outputs[-1], p, f'{prefix}.fusion_blocks.{i}', quant=quant))
    return outputs if complete_output else outputs[-1]


def mv_embedding(points: Tensor, dirs: Tensor, feats: Tensor, p: Dict[str, Tensor], prefix: str,
                 n_views: int, complete_output: bool = False, quant: Optional[str] = None):
    """``MVResNetMLPNeRFEmbedding.call`` (``layers.py:354-379``).

    points/dirs [(b nv), nr, S, 3], feats [(b nv), nr, S, 259]; input order
    [posenc(points) 60 | posenc(dirs) 60 | feats 259] (``layers.py:362``)."""
    enc_p = position_encoding(points)
    enc_d = position_encoding(dirs)
    x = torch.cat([enc_p, enc_d, feats], dim=-1)
    return mv_embedding_from_x(x, p, prefix, n_views, complete_output, quant)


def render_readout(x: Tensor, p: Dict[str, Tensor], prefix: str) -> Tuple[Tensor, Tensor]:
    """``RenderReadout.call`` (``layers.py:392-397``)."""
    y = F.relu(x) @ p[f'{prefix}.output_layer.kernel'] + p[f'{prefix}.output_layer.bias']
    return torch.sigmoid(y[..., :3]), F.softplus(y[..., 3])


# --------------------------------------------------------------------------
# MVVNeRFRenderer._call
# --------------------------------------------------------------------------


def _render_pass(ray_o, ray_d, zs, norm_images, feats, K, Einv, p, net, n_views, quant=None):
    b, nr, S = zs.shape
    pts = ray_o[:, :, None, :] + zs[..., None] * ray_d[:, :, None, :]
    pix, cam = compute_pixel_in_image_mv(pts, K, Einv)
    g = get_projection_features_mv(norm_images, feats, pix)
    cam_dirs = world_to_camera_direction_vector_mv(ray_d, Einv)            # [b, nv, nr, 3]
    rep_dirs = cam_dirs[:, :, :, None, :].expand(-1, -1, -1, S, -1).reshape(b * n_views, nr, S, 3)
    cam3 = cam.reshape(b * n_views, nr, S, 4)[..., :3]
    g = g.reshape(b * n_views, nr, S, -1)
    emb = mv_embedding(cam3, rep_dirs, g, p, f'{net}_embedding', n_views, quant=quant)
    chroma, density = render_readout(emb, p, f'{net}_readout')
    rgb, depth, weights = volumetric_render(zs, density, chroma)
    return rgb, depth, weights, chroma, density


def mvnerf_call(inputs: Sequence[Tensor], combined_features: Tensor, p: Dict[str, Tensor], *,
                near: float, far: float, n_samples: int, u_strat: Tensor, u_pdf: Tensor,
                return_aux: bool = False, quant: Optional[str] = None, feat_up: int = 1):
    """``MVVNeRFRenderer._call`` (``model_v0.py:113-184``).

    ``feat_up == 2``: ``combined_features`` is the [b,nv,H/2,W/2,256] input of the encoder's final
    ``UpSampling2D`` (``layers.py:589``; ``model_without.py:76-79``) and is upsampled here first, the way
    the reference does it -- the two-stage form the CUDA low-resolution gather is tested against.

    inputs = (ray_o [b,nr,3], ray_d [b,nr,3], images [b,nv,H,W,3] in [0,1],
    K [b,nv,4,4], E^-1 [b,nv,4,4]); combined_features [b,nv,H,W,256].
    No ``stop_gradient`` anywhere (quirk Q4).  ``quant``: see ``_quantizer`` (None = reference).
    Returns (rgb, depth, fine_rgb, fine_depth).
    """
    ray_o, ray_d, images, K, Einv = inputs
    n_views = images.shape[1]
    norm_images = images * 2 - 1.0
    if feat_up == 2:
        combined_features = upsample2x_bilinear(combined_features)
    else:
        assert feat_up == 1
    _, z = sample_along_ray(ray_o, ray_d, near, far, n_samples, u_strat)
    rgb, depth, weights, _, _ = _render_pass(ray_o, ray_d, z, norm_images, combined_features,
                                             K, Einv, p, 'coarse', n_views, quant)
    z_mid = 0.5 * (z[..., 1:] + z[..., :-1])
    probs = weights[..., 1:-1]
    z_samples, above, below = sample_pdf(z_mid, probs, n_samples, u_pdf, return_indices=True)
    all_zs = torch.cat([z, z_samples], dim=-1)
    all_zs, perm = torch.sort(all_zs, dim=-1, stable=True)
    fine_rgb, fine_depth, fine_w, _, _ = _render_pass(ray_o, ray_d, all_zs, norm_images, combined_features,
                                                      K, Einv, p, 'fine', n_views, quant)
    if return_aux:
        aux = dict(z=z, weights=weights, z_samples=z_samples, above=above, below=below,
                   all_zs=all_zs, perm=perm.to(torch.int32), fine_weights=fine_w)
        return (rgb, depth, fine_rgb, fine_depth), aux
    return rgb, depth, fine_rgb, fine_depth


def mse(labels: Tensor, pred: Tensor) -> Tensor:
    """``tf.keras.losses.MeanSquaredError`` with default reduction: mean over every element."""
    return torch.mean((pred - labels) ** 2)


def train_loss(outputs, labels: Tensor) -> Tensor:
    """``model_v0.py:193``: MSE(labels, rgb) + MSE(labels, fine_rgb)."""
    rgb, _, fine_rgb, _ = outputs
    return mse(labels, rgb) + mse(labels, fine_rgb)


def clip_gradients(grads: Sequence[Tensor], clip: float = 1.0) -> List[Tensor]:
    """``nerf_utils.py:8-12``: per-element ``clip_by_value``."""
    return [torch.clamp(g, -clip, clip) for g in grads]


def warmup_lr(step: float, target_lr: float, warmup_steps: float, scale_down_after: float = 400000.0) -> float:
    """``WarmupScheduler.__call__`` (``nerf_utils.py:288-300``)."""
    warmup_steps = max(1.0, float(warmup_steps))
    step = float(step)
    if step <= warmup_steps:
        return step / warmup_steps * target_lr
    if step <= scale_down_after:
        return target_lr
    return 0.1 * target_lr


# --------------------------------------------------------------------------
# Synthetic scenes shared by tests, smoke() and bench.py
# --------------------------------------------------------------------------


def polar_pose(azimuth: float, polar: float, radius: float, t_center=(0.6, 0.0, 0.1)) -> np.ndarray:
    """Camera-to-world pose looking at ``t_center`` from a hemisphere, the way the
    reference's ``Affine.polar`` (``dependencies/manipulation_tasks/.../transform.py:58-75``)
    places cameras: position on a sphere of ``radius`` around ``t_center``, optical
    axis (+z) towards the centre."""
    c = np.asarray(t_center, dtype=np.float64)
    pos = c + radius * np.array([math.sin(polar) * math.cos(azimuth),
                                 math.sin(polar) * math.sin(azimuth),
                                 math.cos(polar)])
    zc = c - pos
    zc /= np.linalg.norm(zc)
    up = np.array([0.0, 0.0, 1.0])
    xc = np.cross(zc, up)
    if np.linalg.norm(xc) == 0:
        xc = np.array([math.cos(azimuth), math.sin(azimuth), 0.0])
    else:
        xc /= np.linalg.norm(xc)
    yc = np.cross(zc, xc)
    pose = np.eye(4)
    pose[:3, 0], pose[:3, 1], pose[:3, 2], pose[:3, 3] = xc, yc, zc, pos
    return pose


def synthetic_scene(seed: int, b: int, nv: int, nr: int, H: int, W: int, C: int = 256,
                    n_samples: int = 64, rays: str = 'random', dtype=torch.float32):
    """Synthetic inputs of SURVEY.md 8(d): U[0,1) images, N(0,1) features, hemisphere
    cameras, rays from the target camera (random (row, col) pixels via
    ``bbox_biased_sample`` or the full row-major grid), U[0,1) u_strat / u_pdf / target."""
    rs = np.random.RandomState(seed)
    g = torch.Generator().manual_seed(seed)
    f = 0.5 * W / math.tan(math.radians(30.0))
    K3 = np.array([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1]], dtype=np.float64)
    ro, rd, Ks, Es, pix_rc = [], [], [], [], []
    for _ in range(b):
        poses = [polar_pose(rs.uniform(0, 2 * math.pi), rs.uniform(0.2, 1.1), rs.uniform(0.7, 0.9))
                 for _ in range(nv + 1)]
        tgt = poses[-1]
        if rays == 'grid':
            o, d = get_rays(W, H, tgt, K3.astype(np.float32))
            o, d = o.reshape(-1, 3)[:nr], d.reshape(-1, 3)[:nr]
            rc = np.stack(np.divmod(np.arange(nr), W), axis=1)
        else:
            rc = bbox_biased_sample(nr, np.array([0, 0, H, W]), H, W, rng=rs)
            o, d = get_specific_rays(rc[:, 1], rc[:, 0], tgt, K3.astype(np.float32))
        cfgs = [dict(pose=pp, intrinsics=K3.reshape(-1)) for pp in poses[:-1]]
        e_invs, ks = zip(*[camera_parameters(c) for c in cfgs])
        ro.append(o); rd.append(d); Ks.append(np.array(ks)); Es.append(np.array(e_invs)); pix_rc.append(rc)
    inputs = (torch.tensor(np.array(ro), dtype=torch.float32).to(dtype),
              torch.tensor(np.array(rd), dtype=torch.float32).to(dtype),
              torch.rand((b, nv, H, W, 3), generator=g).to(dtype),
              torch.tensor(np.array(Ks), dtype=torch.float32).to(dtype),
              torch.tensor(np.array(Es), dtype=torch.float32).to(dtype))
    feats = torch.randn((b, nv, H, W, C), generator=g).to(dtype)
    u_strat = torch.rand((b, nr, n_samples), generator=g).to(dtype)
    u_pdf = torch.rand((b, nr, n_samples), generator=g).to(dtype)
    target = torch.rand((b, nr, 3), generator=g).to(dtype)
    return dict(inputs=inputs, features=feats, u_strat=u_strat, u_pdf=u_pdf, target=target,
                pixels_rc=np.array(pix_rc))


# --------------------------------------------------------------------------
# Pose-query field: DeltaNGF / LanguageNeRF / GraspMVNeRF `_call` + GraspReadout
# (src/lib/delta_ngf/model.py, src/lib/lmvnerf/model_v4.py, src/lib/grasp_mvnerf/model.py)
# --------------------------------------------------------------------------


def _affine_matrix(translation, euler_xyz=None) -> np.ndarray:
    """``manipulation_tasks.transform.Affine(translation, rotation)`` for the two forms the models use
    (dependencies/manipulation_tasks/manipulation_tasks/transform.py:11-23): identity rotation or
    ``Rotation.from_euler('xyz', r)`` (scipy, as the reference)."""
    from scipy.spatial.transform import Rotation
    m = np.eye(4)
    m[:3, 3] = np.array(translation)
    m[:3, :3] = Rotation.from_quat((0, 0, 0, 1)).as_matrix() if euler_xyz is None else \
        Rotation.from_euler('xyz', euler_xyz).as_matrix()
    return m


def transforms_to_check(n_5d_poses: int = 7) -> Tensor:
    """The fixed query offsets of one grasp pose, [6 * n_5d_poses, 4, 4] fp32, base-major
    (``delta_ngf/model.py:63-97``): 6 gripper bases x 7 steps along the local z axis."""
    ox, oy, oz = 0.02, 0.015, 0.0125
    step = (ox - 0.005) / ((n_5d_poses - 1) / 2)
    bases = [_affine_matrix([0, oy, 0]), _affine_matrix([0, -oy, 0]),
             _affine_matrix([-ox, oy, oz], [0.0, np.pi / 2, 0.0]), _affine_matrix([ox, oy, oz], [0.0, -np.pi / 2, 0.0]),
             _affine_matrix([-ox, -oy, oz], [0.0, np.pi / 2, 0.0]), _affine_matrix([ox, -oy, oz], [0.0, -np.pi / 2, 0.0])]
    c = int((n_5d_poses - 1) / 2)
    steps = [_affine_matrix([0.0, 0.0, i * step]) for i in range(-c, c + 1)]
    return torch.tensor(np.array([b @ t for b in bases for t in steps]), dtype=torch.float32)


def rotation_from_quaternion(q: Tensor) -> Tensor:
    """``tensorflow_graphics.geometry.transformation.rotation_matrix_3d.from_quaternion`` (third party,
    unpinned version; call site ``delta_ngf/model.py:29``): q = [x, y, z, w], assumed normalised."""
    x, y, z, w = q.unbind(-1)
    tx, ty, tz = 2.0 * x, 2.0 * y, 2.0 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    m = torch.stack([1.0 - (tyy + tzz), txy - twz, txz + twy,
                     txy + twz, 1.0 - (txx + tzz), tyz - twx,
                     txz - twy, tyz + twx, 1.0 - (txx + tyy)], dim=-1)
    return m.reshape(*q.shape[:-1], 3, 3)


def t_m_to_h_matrix(translations: Tensor, rot: Tensor) -> Tensor:
    """``delta_ngf/model.py:16-25``: [[R | t], [0 0 0 1]]."""
    top = torch.cat([rot, translations[..., None]], dim=-1)
    last = torch.tensor([0., 0., 0., 1.], dtype=top.dtype).expand(*top.shape[:-2], 1, 4)
    return torch.cat([top, last], dim=-2)


def compute_matrices(translations: Tensor, rotations: Tensor, rotation_representation: str = 'quaternion') -> Tensor:
    """``DeltaNGF.compute_matrices`` (``delta_ngf/model.py:172-183``): the 6-d form normalises both vectors
    WITHOUT orthogonalising the second one."""
    if rotation_representation == 'quaternion':
        return t_m_to_h_matrix(translations, rotation_from_quaternion(rotations))
    if rotation_representation == '6d':
        r1 = rotations[..., :3] / rotations[..., :3].norm(dim=-1, keepdim=True)
        r2 = rotations[..., 3:] / rotations[..., 3:].norm(dim=-1, keepdim=True)
        r3 = torch.linalg.cross(r1, r2, dim=-1)
        return t_m_to_h_matrix(translations, torch.stack([r1, r2, r3], dim=-1))
    raise ValueError('Unknown rotation representation: ' + rotation_representation)


def grasp_readout_param_shapes(use_bias: bool = True):
    """``GraspReadout`` variables (``delta_ngf/layers.py:8-28``): 582 337 parameters with use_bias."""
    s = []
    for i in range(1, 5):
        s += [(f'activation_downscale_{i}.kernel', (128, 64)), (f'activation_downscale_{i}.bias', (64,))]
    s += [('combined_activation_downscale.kernel', (256, 64)), ('combined_activation_downscale.bias', (64,))]
    s += [('readout.0.layer_0.kernel', (2688, 128)), ('readout.0.layer_0.bias', (128,)),
          ('readout.0.layer_1.kernel', (128, 64)), ('readout.0.layer_1.bias', (64,)),
          ('readout.0.shortcut.kernel', (2688, 64)),
          ('readout.1.layer_0.kernel', (64, 64)), ('readout.1.layer_0.bias', (64,)),
          ('readout.1.layer_1.kernel', (64, 64)), ('readout.1.layer_1.bias', (64,)),
          ('readout.2.output_layer.kernel', (64, 1))]
    if use_bias:
        s.append(('readout.2.output_layer.bias', (1,)))
    return s


def init_grasp_readout(seed: int, bias_std: float = 0.0, use_bias: bool = True, dtype=torch.float32) -> Dict[str, Tensor]:
    """he_normal kernels (``delta_ngf/layers.py``; glorot for the combined downscale), zero biases."""
    g = torch.Generator().manual_seed(seed)
    p = {}
    for name, shape in grasp_readout_param_shapes(use_bias):
        full = f'grasp_readout.{name}'
        if name.endswith('kernel'):
            p[full] = (torch.randn(shape, generator=g) * math.sqrt(2.0 / shape[0])).to(dtype)
        else:
            p[full] = (torch.randn(shape, generator=g) * bias_std).to(dtype)
    return p


def grasp_readout(embeddings: Sequence[Tensor], p: Dict[str, Tensor], prefix: str = 'grasp_readout',
                  activation: str = 'elu', quant: Optional[str] = None, return_pre_relu: bool = False):
    """``GraspReadout.call`` (``delta_ngf/layers.py:30-42``; relu variant ``grasp_mvnerf/layers.py:22-31``):
    4 x [b, np, n5, 128] -> [b, np] (``return_pre_relu``: also the input of the final Readout's ReLU, [b, np, 64]).  ``quant='bf16'`` rounds every matmul operand (inputs, kernels, activation
    outputs) where the tcgen05 head (``cn_grasp_head_fwd``) stores bf16 operand tiles."""
    q = _quantizer(quant)
    act = F.elu if activation == 'elu' else F.relu
    ds = [q(act(q(embeddings[i]) @ q(p[f'{prefix}.activation_downscale_{i + 1}.kernel'])
                + p[f'{prefix}.activation_downscale_{i + 1}.bias'])) for i in range(4)]
    c = torch.cat(ds, dim=-1)
    c = q(act(c @ q(p[f'{prefix}.combined_activation_downscale.kernel']) + p[f'{prefix}.combined_activation_downscale.bias']))
    c = c.reshape(*c.shape[:-2], -1)                                   # 'b np n5 d -> b np (n5 d)'
    y = resnet_mlp_block(c, p, f'{prefix}.readout.0', activation=activation,
                         shortcut_kernel=q(p[f'{prefix}.readout.0.shortcut.kernel']), quant=quant)
    o2 = resnet_mlp_block(y, p, f'{prefix}.readout.1', activation=activation, quant=quant)
    y = F.relu(o2) @ p[f'{prefix}.readout.2.output_layer.kernel']
    if f'{prefix}.readout.2.output_layer.bias' in p:
        y = y + p[f'{prefix}.readout.2.output_layer.bias']
    return (y[..., 0], o2) if return_pre_relu else y[..., 0]


def query_embeddings_from_points(images: Tensor, K: Tensor, Einv: Tensor, pts: Tensor, dirs: Tensor, features: Tensor,
                                 p: Dict[str, Tensor], net: str = 'fine', quant: Optional[str] = None) -> List[Tensor]:
    """the part of ``DeltaNGF._call`` between the query points and the ``GraspReadout``
    (``delta_ngf/model.py:202-241``): world points / approach directions [b, n, 3] -> the last four
    `complete_output` stage outputs of the fine embedding, each [b, n, 128]."""
    b, nv = images.shape[:2]
    n = pts.shape[1]
    pix, cam = compute_pixel_in_image_mv(pts[:, :, None, :], K, Einv)    # [b, nv, n, 1, 2], [..., 4]
    g = get_projection_features_mv(images * 2 - 1.0, features, pix)
    cam_dirs = world_to_camera_direction_vector_mv(dirs, Einv)           # [b, nv, n, 3]
    cam3 = cam.reshape(b * nv, n, 1, 4)[..., :3]
    cam_dirs = cam_dirs.reshape(b * nv, n, 1, 3)
    g = g.reshape(b * nv, n, 1, -1)
    out = mv_embedding(cam3, cam_dirs, g, p, f'{net}_embedding', nv, complete_output=True, quant=quant)[4:]
    return [o[:, :, 0] for o in out]


def query_embeddings(images: Tensor, K: Tensor, Einv: Tensor, transforms: Tensor, features: Tensor,
                     p: Dict[str, Tensor], net: str = 'fine', quant: Optional[str] = None,
                     n_5d_poses: int = 7) -> List[Tensor]:
    """``DeltaNGF._call`` up to the embeddings (``delta_ngf/model.py:188-241``): transforms [b, np, 4, 4] ->
    the last four `complete_output` stage outputs, each [b, np, n5, 128]."""
    b = images.shape[0]
    n_pts = transforms.shape[1]
    ttc = transforms_to_check(n_5d_poses).to(transforms.dtype)
    n5 = ttc.shape[0]
    poses = transforms[:, :, None] @ ttc[None, None]                   # [b, np, n5, 4, 4] (row order np, n5)
    pts = poses[..., :3, 3].reshape(b, n_pts * n5, 3)
    dirs = poses[..., :3, 2].reshape(b, n_pts * n5, 3)                 # R . z_dir
    out = query_embeddings_from_points(images, K, Einv, pts, dirs, features, p, net, quant)
    return [o.reshape(b, n_pts, n5, -1) for o in out]


def query_call(images: Tensor, K: Tensor, Einv: Tensor, transforms: Tensor, features: Tensor,
               p: Dict[str, Tensor], net: str = 'fine', quant: Optional[str] = None, activation: str = 'elu',
               head_quant: Optional[str] = None, return_pre_relu: bool = False):
    """``DeltaNGF._call`` (``delta_ngf/model.py:188-245``) -> grasp_success [b, np].  ``quant`` rounds the
    embedding's operands, ``head_quant`` the ``GraspReadout``'s (the tcgen05 head; the torch head runs TF32)."""
    return grasp_readout(query_embeddings(images, K, Einv, transforms, features, p, net, quant), p,
                         activation=activation, quant=head_quant, return_pre_relu=return_pre_relu)


def cosine_similarity_loss(y_true: Tensor, y_pred: Tensor) -> Tensor:
    """``tf.keras.losses.CosineSimilarity(axis=-1)``: -mean(sum(l2norm(y_true) * l2norm(y_pred)))."""
    def l2n(x):
        return x * torch.rsqrt(torch.clamp((x * x).sum(-1, keepdim=True), min=1e-12))
    return -(l2n(y_true) * l2n(y_pred)).sum(-1).mean()


def synthetic_poses(seed: int, b: int, n_pts: int, representation: str = 'quaternion', dtype=torch.float32):
    """grasp poses around the scene centre the synthetic cameras look at: translations, rotations"""
    g = torch.Generator().manual_seed(seed + 17)
    t = torch.tensor([0.6, 0.0, 0.1]) + (torch.rand((b, n_pts, 3), generator=g) - 0.5) * 0.1
    if representation == 'quaternion':
        q = torch.randn((b, n_pts, 4), generator=g)
        r = q / q.norm(dim=-1, keepdim=True)
    else:
        r = torch.randn((b, n_pts, 6), generator=g)
    return t.to(dtype), r.to(dtype)


def delta_ngf_losses(images, K, Einv, features, p, landscape_pose, trajectory_pose, labels, *,
                     rotation_representation: str = 'quaternion', quant: Optional[str] = None,
                     loss_fn=None) -> Dict[str, Tensor]:
    """the loss terms of ``DeltaNGF.train_step`` (``delta_ngf/model.py:258-291``): landscape loss at
    ``landscape_pose = (T, R)`` (inputs[0:2]) + cosine similarity between d(sum prediction)/d(pose variables) at
    ``trajectory_pose`` (inputs[2:4]) and ``labels[1]`` / ``labels[2]``.  Differentiable w.r.t. the
    ``grasp_readout.*`` entries of ``p`` (the only variables the reference trains)."""
    loss_fn = mse if loss_fn is None else loss_fn
    y_pred = query_call(images, K, Einv, compute_matrices(*landscape_pose, rotation_representation), features, p,
                        quant=quant)
    landscape = loss_fn(labels[0], y_pred)
    t = trajectory_pose[0].detach().clone().requires_grad_(True)
    r = trajectory_pose[1].detach().clone().requires_grad_(True)
    pred = query_call(images, K, Einv, compute_matrices(t, r, rotation_representation), features, p, quant=quant)
    g_t, g_r = torch.autograd.grad(pred.sum(), (t, r), create_graph=True)
    loss_t = cosine_similarity_loss(labels[1], g_t)
    if rotation_representation == 'quaternion':
        loss_r = cosine_similarity_loss(labels[2], g_r)
    else:
        loss_r = cosine_similarity_loss(labels[2][..., :3], g_r[..., :3]) + \
            cosine_similarity_loss(labels[2][..., 3:], g_r[..., 3:])
    return {'loss': loss_t + loss_r + landscape, 'landscape_loss': landscape, 'grad_loss_t': loss_t,
            'grad_loss_r': loss_r, 'prediction': pred, 'output_gradients': (g_t, g_r)}
