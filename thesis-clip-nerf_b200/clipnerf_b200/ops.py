"""One thin torch wrapper per C-ABI entry point (``include/clipnerf_b200.h``).

Every function takes CUDA tensors, allocates its outputs with torch (the library owns no device
memory), enqueues on torch's current stream and returns without synchronising."""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from . import _abi
from ._abi import ACT_TILE_BYTES, PACKED_BYTES, SLAB_BYTES, TILE_ROWS, X_TILE_BYTES, CnScene, call


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32c(t: torch.Tensor) -> torch.Tensor:
    assert t.is_cuda, 'clipnerf_b200 ops need CUDA tensors (no CPU fallback)'
    if t.dtype != torch.float32 or not t.is_contiguous():
        t = t.to(torch.float32).contiguous()
    return t


def _bytes(n: int, device) -> torch.Tensor:
    return torch.empty(n, dtype=torch.uint8, device=device)


class Scene:
    """Source views: images [b,nv,H,W,3], features [b,nv,H,W,256], K [b,nv,4,4], E^-1 [b,nv,4,4].

    features [b,nv,H/2,W/2,256] = the input of the encoder's final 2x bilinear UpSampling2D (layers.py:589):
    the upsample is folded into the gather (CnScene.feat_up = 2) and every feature gradient is low-resolution."""

    def __init__(self, images, features, intrinsics, extr_inv):
        self.images, self.features = _f32c(images), _f32c(features)
        self.intrinsics, self.extr_inv = _f32c(intrinsics), _f32c(extr_inv)
        b, nv, H, W, c3 = self.images.shape
        shape = tuple(self.features.shape)
        if shape == (b, nv, H, W, 256):
            self.feat_up = 1
        else:
            assert c3 == 3 and H % 2 == 0 and W % 2 == 0 and shape == (b, nv, H // 2, W // 2, 256), \
                'features must be [b,nv,H,W,256] or the pre-upsample [b,nv,H/2,W/2,256]'
            self.feat_up = 2
        assert c3 == 3
        assert tuple(self.intrinsics.shape) == (b, nv, 4, 4) and tuple(self.extr_inv.shape) == (b, nv, 4, 4)
        self.b, self.nv, self.H, self.W = b, nv, H, W
        self.c = CnScene(self.images.data_ptr(), self.features.data_ptr(), self.intrinsics.data_ptr(),
                         self.extr_inv.data_ptr(), b, nv, H, W, self.feat_up)

    @property
    def device(self):
        return self.images.device


def sample_coarse(u: torch.Tensor, near: float, far: float) -> torch.Tensor:
    u = _f32c(u)
    z = torch.empty_like(u)
    call('cn_sample_coarse', _ptr(u), _ptr(z), u.numel() // u.shape[-1], u.shape[-1], float(near), float(far),
         _stream())
    return z


def sample_pdf_fwd(z_c, w_c, u) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    z_c, w_c, u = _f32c(z_c), _f32c(w_c), _f32c(u)
    assert z_c.shape[-1] == 64 and w_c.shape == z_c.shape and u.shape == z_c.shape
    n = z_c.numel() // 64
    z_fine = torch.empty(*z_c.shape[:-1], 128, dtype=torch.float32, device=z_c.device)
    perm = torch.empty(*z_c.shape[:-1], 128, dtype=torch.uint8, device=z_c.device)
    above = torch.empty(*z_c.shape[:-1], 64, dtype=torch.uint8, device=z_c.device)
    call('cn_sample_pdf_fwd', _ptr(z_c), _ptr(w_c), _ptr(u), n, _ptr(z_fine), _ptr(perm), _ptr(above), _stream())
    return z_fine, perm, above


def sample_pdf_bwd(z_c, w_c, u, perm, dz_fine, dw_c: Optional[torch.Tensor] = None) -> torch.Tensor:
    z_c, w_c, u, dz_fine = _f32c(z_c), _f32c(w_c), _f32c(u), _f32c(dz_fine)
    if dw_c is None:
        dw_c = torch.zeros_like(w_c)
    call('cn_sample_pdf_bwd', _ptr(z_c), _ptr(w_c), _ptr(u), _ptr(perm), _ptr(dz_fine), z_c.numel() // 64,
         _ptr(dw_c), _stream())
    return dw_c


def n_tiles(rows: int) -> int:
    assert rows % TILE_ROWS == 0, f'{rows} rows is not a multiple of {TILE_ROWS}'
    return rows // TILE_ROWS


def gather_workspace(scene: Scene, nr: int, S: int) -> torch.Tensor:
    """caller-owned workspace of the binned gather (sorted row records; filled by the forward or by
    ``project_bin`` and re-used by the backward of the same scene / rays / z)."""
    n = _abi.load().cn_gather_ws_bytes(scene.b, scene.nv, scene.H, scene.W, nr, S)
    assert n > 0
    return _bytes(n, scene.device)


def project_bin(scene: Scene, ray_o, ray_d, z, ws: Optional[torch.Tensor] = None) -> torch.Tensor:
    ray_o, ray_d, z = _f32c(ray_o), _f32c(ray_d), _f32c(z)
    b, nr, S = z.shape
    if ws is None:
        ws = gather_workspace(scene, nr, S)
    call('cn_project_bin', ctypes.byref(scene.c), _ptr(ray_o), _ptr(ray_d), _ptr(z), nr, S, _ptr(ws), ws.numel(),
         _stream())
    return ws


def project_gather_fwd(scene: Scene, ray_o, ray_d, z, ws: Optional[torch.Tensor] = None, return_ws: bool = False):
    ray_o, ray_d, z = _f32c(ray_o), _f32c(ray_d), _f32c(z)
    b, nr, S = z.shape
    if ws is None:
        ws = gather_workspace(scene, nr, S)
    x = _bytes(n_tiles(b * scene.nv * nr * S) * X_TILE_BYTES, z.device)
    call('cn_project_gather_fwd', ctypes.byref(scene.c), _ptr(ray_o), _ptr(ray_d), _ptr(z), nr, S, _ptr(x), _ptr(ws),
         ws.numel(), _stream())
    return (x, ws) if return_ws else x


def project_gather_debug(scene: Scene, ray_o, ray_d, z, gather: bool = True, cam: bool = False):
    ray_o, ray_d, z = _f32c(ray_o), _f32c(ray_d), _f32c(z)
    b, nr, S = z.shape
    dev = z.device
    pix = torch.empty(b, scene.nv, nr, S, 2, dtype=torch.float32, device=dev)
    corner = torch.empty(b, scene.nv, nr, S, 2, dtype=torch.int32, device=dev)
    g = torch.empty(b, scene.nv, nr, S, 259, dtype=torch.float32, device=dev) if gather else None
    c = torch.empty(b, scene.nv, nr, S, 4, dtype=torch.float32, device=dev) if cam else None
    call('cn_project_gather_debug', ctypes.byref(scene.c), _ptr(ray_o), _ptr(ray_d), _ptr(z), nr, S, _ptr(pix),
         _ptr(corner), _ptr(g), _ptr(c), _stream())
    return (pix, corner, g, c) if cam else (pix, corner, g)


def interpolate_bilinear(grid, query):
    grid, query = _f32c(grid), _f32c(query)
    B, H, W, C = grid.shape
    N = query.shape[1]
    out = torch.empty(B, N, C, dtype=torch.float32, device=grid.device)
    call('cn_interpolate_bilinear', _ptr(grid), _ptr(query), B, H, W, C, N, _ptr(out), _stream())
    return out


def position_encoding(pos, n_freq: int = 10, freq0: float = 3.14159274101257324):
    pos = _f32c(pos)
    D = pos.shape[-1]
    out = torch.empty(*pos.shape[:-1], D * n_freq * 2, dtype=torch.float32, device=pos.device)
    call('cn_position_encoding', _ptr(pos), pos.numel() // D, D, n_freq, float(freq0), _ptr(out), _stream())
    return out


def sample_pdf(bins, weights, u, return_above: bool = False):
    bins, weights, u = _f32c(bins), _f32c(weights), _f32c(u)
    assert bins.shape[-1] == 63 and weights.shape[-1] == 62 and u.shape[-1] == 64, 'sample_pdf is built for 64 samples'
    samples = torch.empty_like(u)
    above = torch.empty(u.shape, dtype=torch.int32, device=u.device) if return_above else None
    call('cn_sample_pdf', _ptr(bins), _ptr(weights), _ptr(u), u.numel() // 64, _ptr(samples), _ptr(above), _stream())
    return (samples, above) if return_above else samples


def project_gather_bwd(scene: Scene, ray_o, ray_d, z, dx_tiles, d_features: Optional[torch.Tensor],
                       dz: Optional[torch.Tensor], ws: Optional[torch.Tensor] = None) -> None:
    """``ws``: the workspace the forward filled for the same (scene, rays, z); rebuilt when omitted."""
    ray_o, ray_d, z = _f32c(ray_o), _f32c(ray_d), _f32c(z)
    b, nr, S = z.shape
    if ws is None:
        ws = project_bin(scene, ray_o, ray_d, z)
    n_kernels = (2 if dz is not None else 0) + (2 if d_features is not None else 0)
    call('cn_project_gather_bwd', ctypes.byref(scene.c), _ptr(ray_o), _ptr(ray_d), _ptr(z), nr, S, _ptr(dx_tiles),
         _ptr(ws), ws.numel(), _ptr(d_features), _ptr(dz), _stream(), extra_launches=n_kernels)


def project_gather_bwd_rays(scene: Scene, ray_o, ray_d, z, dx_tiles, ws: torch.Tensor, need_dz: bool = False):
    """(d_ray_o [b,nr,3], d_ray_d [b,nr,3], dz [b,nr,S] or None) of the rows whose MLP-input gradient is ``dx_tiles``."""
    ray_o, ray_d, z = _f32c(ray_o), _f32c(ray_d), _f32c(z)
    b, nr, S = z.shape
    d_o, d_d = torch.zeros_like(ray_o), torch.zeros_like(ray_d)
    dz = torch.zeros_like(z) if need_dz else None
    call('cn_project_gather_bwd_rays', ctypes.byref(scene.c), _ptr(ray_o), _ptr(ray_d), _ptr(z), nr, S, _ptr(dx_tiles),
         _ptr(ws), ws.numel(), _ptr(dz), _ptr(d_o), _ptr(d_d), _stream())
    return d_o, d_d, dz


def feature_grad(scene: Scene, sets, d_features: torch.Tensor, accumulate: bool) -> None:
    """d_features (+)= feature-map gradient of every (ws, dx_tiles, nr, S) in ``sets`` in one sweep."""
    n = len(sets)
    assert 1 <= n <= 4
    I32, SZ, VP = ctypes.c_int32 * n, ctypes.c_size_t * n, ctypes.c_void_p * n
    nr = I32(*[s[2] for s in sets])
    S = I32(*[s[3] for s in sets])
    ws = VP(*[s[0].data_ptr() for s in sets])
    wsb = SZ(*[s[0].numel() for s in sets])
    dx = VP(*[s[1].data_ptr() for s in sets])
    assert d_features.is_contiguous() and d_features.dtype == torch.float32
    call('cn_feature_grad', ctypes.byref(scene.c), n, nr, S, ws, wsb, dx, _ptr(d_features), int(accumulate), _stream())


def generate_rays(pose, intrinsics, W: int, first_pixel: int, n: int, device):
    """rays of pixels [first_pixel, first_pixel + n) of the target view (row-major), generated on the device"""
    import numpy as np
    pose, intrinsics = np.asarray(pose), np.asarray(intrinsics)
    m = pose[:3, :3] @ np.linalg.inv(intrinsics[:3, :3])             # the reference's own host expression
    single = int(m.dtype == np.float32)
    m64 = np.ascontiguousarray(m, dtype=np.float64)
    o32 = np.ascontiguousarray(pose[:3, 3], dtype=np.float32)
    ray_o = torch.empty(n, 3, dtype=torch.float32, device=device)
    ray_d = torch.empty(n, 3, dtype=torch.float32, device=device)
    call('cn_generate_rays', m64.ctypes.data_as(ctypes.c_void_p), o32.ctypes.data_as(ctypes.c_void_p), single, int(W),
         int(first_pixel), int(n), _ptr(ray_o), _ptr(ray_d), _stream())
    return ray_o, ray_d


def quantize_render(rgb, depth):
    """(uint8 [n,3], uint8 [n]) of model_v0.py:272-280: x255 clip truncate; depth min-max normalised, x255, truncate"""
    rgb, depth = _f32c(rgb), _f32c(depth)
    n = depth.numel()
    assert rgb.numel() == 3 * n
    ws = torch.empty(2, dtype=torch.int32, device=rgb.device)
    rgb8 = torch.empty(n, 3, dtype=torch.uint8, device=rgb.device)
    depth8 = torch.empty(n, dtype=torch.uint8, device=rgb.device)
    call('cn_quantize_render', _ptr(rgb), _ptr(depth), n, _ptr(ws), _ptr(rgb8), _ptr(depth8), _stream())
    return rgb8, depth8


def images_u8_to_f32(src: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """uint8 [..., H, W, C>=3] -> float32 [..., H, W, 3] = float32(double(u8) / 255.0), the reference's host expression"""
    assert src.is_cuda and src.dtype == torch.uint8 and src.is_contiguous() and src.shape[-1] >= 3
    if out is None:
        out = torch.empty(*src.shape[:-1], 3, dtype=torch.float32, device=src.device)
    assert out.is_contiguous() and out.dtype == torch.float32 and out.numel() == src.numel() // src.shape[-1] * 3
    call('cn_images_u8_to_f32', _ptr(src), src.numel() // src.shape[-1], src.shape[-1], _ptr(out), _stream())
    return out


def composite_fwd(z, rgbs):
    z, rgbs = _f32c(z), _f32c(rgbs)
    S = z.shape[-1]
    n = z.numel() // S
    dev = z.device
    rgb = torch.empty(*z.shape[:-1], 3, dtype=torch.float32, device=dev)
    depth = torch.empty(*z.shape[:-1], dtype=torch.float32, device=dev)
    w = torch.empty_like(z)
    call('cn_composite_fwd', _ptr(z), _ptr(rgbs), n, S, _ptr(rgb), _ptr(depth), _ptr(w), _stream())
    return rgb, depth, w


def composite_bwd(z, rgbs, d_rgb, d_depth=None, d_weights=None, need_dz=False):
    z, rgbs, d_rgb = _f32c(z), _f32c(rgbs), _f32c(d_rgb)
    d_depth = None if d_depth is None else _f32c(d_depth)
    d_weights = None if d_weights is None else _f32c(d_weights)
    S = z.shape[-1]
    n = z.numel() // S
    d_y = torch.empty(*z.shape, 4, dtype=torch.float32, device=z.device)
    dz = torch.empty_like(z) if need_dz else None
    call('cn_composite_bwd', _ptr(z), _ptr(rgbs), n, S, _ptr(d_rgb), _ptr(d_depth), _ptr(d_weights), _ptr(d_y),
         _ptr(dz), _stream())
    return d_y, dz


# activation tiles saved per 128-row tile for the backward: relu(h) of the 3 blocks of a stage (the hidden
# activation relu(t) is recomputed by cn_mlp_block_bwd); the fusion stage also keeps relu(e) for the readout
FEATURE_ACTS = 3
FUSION_ACTS = 4


def pack_weights(params: torch.Tensor) -> torch.Tensor:
    params = _f32c(params)
    assert params.numel() == _abi.NET_PARAMS
    packed = _bytes(PACKED_BYTES, params.device)
    call('cn_pack_weights', _ptr(params), _ptr(packed), _stream())
    return packed


def mlp_feature_fwd(packed, x_tiles, save_acts: bool, save_hidden: bool = False, mask_acts=None):
    """per-view stage.  ``save_hidden``: acts holds FEATURE_ACTS + 3 tiles per tile (the hidden activations behind the
    usual ones); ``mask_acts``: the tangent pass with the ReLU patterns of that buffer (nothing is saved)."""
    nt = x_tiles.numel() // X_TILE_BYTES
    h3 = _bytes(nt * ACT_TILE_BYTES, x_tiles.device)
    if not save_hidden and mask_acts is None:
        acts = _bytes(nt * FEATURE_ACTS * ACT_TILE_BYTES, x_tiles.device) if save_acts else None
        call('cn_mlp_feature_fwd', _ptr(packed), _ptr(x_tiles), nt, _ptr(h3), _ptr(acts), _stream())
        return h3, acts
    acts = _bytes(nt * (FEATURE_ACTS + 3) * ACT_TILE_BYTES, x_tiles.device) if (save_acts and mask_acts is None) else None
    call('cn_mlp_feature_fwd_ex', _ptr(packed), _ptr(x_tiles), nt, _ptr(h3), _ptr(acts), int(save_hidden), _ptr(mask_acts),
         _stream())
    return h3, acts


def query_tangent_x(scene: Scene, pts, drs, u_pts, u_dirs) -> torch.Tensor:
    """tangent of the MLP-input rows of the pose-query call: pts, drs, u_* [b, n, 3] (n % 128 == 0) -> X tiles"""
    pts, drs, u_pts, u_dirs = _f32c(pts), _f32c(drs), _f32c(u_pts), _f32c(u_dirs)
    b, n, _ = pts.shape
    x = _bytes(n_tiles(b * scene.nv * n) * X_TILE_BYTES, pts.device)
    call('cn_query_tangent_x', ctypes.byref(scene.c), _ptr(pts), _ptr(drs), _ptr(u_pts), _ptr(u_dirs), n, _ptr(x), _stream())
    return x


def mlp_chain_fwd(packed, x_tiles, save_acts: bool, taps: Optional[str] = None, save_hidden: bool = False,
                  mask_acts_f=None, mask_acts_u=None):
    """ONE source view: per-view stage, fusion stage and readout in one launch (bit-identical to the two-kernel path).
    ``taps``: None, 'f32' (rows [4, rows, 128]) or 'bf16' (tile images [4, n_tiles * ACT_TILE_BYTES]).
    Returns (rgbs [rows, 4], acts_f, acts_u, taps)."""
    nt = x_tiles.numel() // X_TILE_BYTES
    dev = x_tiles.device
    rows = nt * TILE_ROWS
    rgbs = torch.empty(rows, 4, dtype=torch.float32, device=dev)
    tangent = mask_acts_f is not None
    extra = 3 if (save_hidden or tangent) else 0
    acts_f = _bytes(nt * (FEATURE_ACTS + extra) * ACT_TILE_BYTES, dev) if (save_acts and not tangent) else None
    acts_u = _bytes(nt * (FUSION_ACTS + extra) * ACT_TILE_BYTES, dev) if (save_acts and not tangent) else None
    t, mode = None, 0
    if taps == 'f32':
        t, mode = torch.empty(4, rows, 128, dtype=torch.float32, device=dev), 1
    elif taps == 'bf16':
        t, mode = torch.empty(4, nt * ACT_TILE_BYTES, dtype=torch.uint8, device=dev), 2
    call('cn_mlp_chain_fwd', _ptr(packed), _ptr(x_tiles), nt, _ptr(rgbs), _ptr(acts_f), _ptr(acts_u), _ptr(t), mode,
         int(save_hidden), _ptr(mask_acts_f), _ptr(mask_acts_u), _stream())
    return rgbs, acts_f, acts_u, t


def mlp_fusion_fwd(packed, h3_tiles, b: int, nv: int, save_acts: bool):
    nt = h3_tiles.numel() // ACT_TILE_BYTES
    assert nt % (b * nv) == 0
    tpv = nt // (b * nv)
    rgbs = torch.empty(b * tpv * TILE_ROWS, 4, dtype=torch.float32, device=h3_tiles.device)
    acts = _bytes(b * tpv * FUSION_ACTS * ACT_TILE_BYTES, h3_tiles.device) if save_acts else None
    call('cn_mlp_fusion_fwd', _ptr(packed), _ptr(h3_tiles), b, nv, tpv, _ptr(rgbs), _ptr(acts), _stream())
    return rgbs, acts


def mlp_fusion_fwd_taps(packed, h3_tiles, b: int, nv: int, save_acts: bool, save_hidden: bool = False, mask_acts=None):
    """fusion stage returning the four `complete_output` stage outputs: taps [4, b*tpv*128, 128] fp32.
    ``save_hidden`` / ``mask_acts`` as for ``mlp_feature_fwd`` (FUSION_ACTS + 3 tiles per tile)."""
    nt = h3_tiles.numel() // ACT_TILE_BYTES
    assert nt % (b * nv) == 0
    tpv = nt // (b * nv)
    rows = b * tpv * TILE_ROWS
    rgbs = torch.empty(rows, 4, dtype=torch.float32, device=h3_tiles.device)
    taps = torch.empty(4, rows, 128, dtype=torch.float32, device=h3_tiles.device)
    if not save_hidden and mask_acts is None:
        acts = _bytes(b * tpv * FUSION_ACTS * ACT_TILE_BYTES, h3_tiles.device) if save_acts else None
        call('cn_mlp_fusion_fwd_taps', _ptr(packed), _ptr(h3_tiles), b, nv, tpv, _ptr(rgbs), _ptr(acts), _ptr(taps),
             _stream())
        return taps, rgbs, acts
    acts = _bytes(b * tpv * (FUSION_ACTS + 3) * ACT_TILE_BYTES, h3_tiles.device) if (save_acts and mask_acts is None) else None
    call('cn_mlp_fusion_fwd_taps_ex', _ptr(packed), _ptr(h3_tiles), b, nv, tpv, _ptr(rgbs), _ptr(acts), _ptr(taps), 0,
         int(save_hidden), _ptr(mask_acts), _stream())
    return taps, rgbs, acts


def mlp_fusion_fwd_tap_tiles(packed, h3_tiles, b: int, nv: int, save_acts: bool):
    """fusion stage returning the four `complete_output` stage outputs as bf16 activation-tile images
    [4, n_tiles * ACT_TILE_BYTES] uint8 (the operand tiles of ``grasp_head_fwd``)."""
    nt = h3_tiles.numel() // ACT_TILE_BYTES
    assert nt % (b * nv) == 0
    tpv = nt // (b * nv)
    rows = b * tpv * TILE_ROWS
    rgbs = torch.empty(rows, 4, dtype=torch.float32, device=h3_tiles.device)
    taps = torch.empty(4, b * tpv * ACT_TILE_BYTES, dtype=torch.uint8, device=h3_tiles.device)
    acts = _bytes(b * tpv * FUSION_ACTS * ACT_TILE_BYTES, h3_tiles.device) if save_acts else None
    call('cn_mlp_fusion_fwd_tap_tiles', _ptr(packed), _ptr(h3_tiles), b, nv, tpv, _ptr(rgbs), _ptr(acts), _ptr(taps),
         _stream())
    return taps, rgbs, acts


def tile_add_rows(g_tiles: Optional[torch.Tensor], rows: torch.Tensor) -> torch.Tensor:
    """g_tiles (bf16 tile images; created when None) (+)= rows [n_tiles*128, 128] fp32"""
    rows = _f32c(rows)
    assert rows.shape[-1] == 128
    nt = n_tiles(rows.numel() // 128)
    acc = g_tiles is not None
    if g_tiles is None:
        g_tiles = _bytes(nt * ACT_TILE_BYTES, rows.device)
    call('cn_tile_add_rows', _ptr(g_tiles), _ptr(rows), nt, int(acc), _stream())
    return g_tiles


def tile_add_tiles(g_tiles: torch.Tensor, src_tiles: torch.Tensor) -> torch.Tensor:
    """g_tiles += src_tiles (both bf16 activation-tile images)"""
    assert g_tiles.numel() == src_tiles.numel() and g_tiles.numel() % ACT_TILE_BYTES == 0
    call('cn_tile_add_tiles', _ptr(g_tiles), _ptr(src_tiles), g_tiles.numel() // ACT_TILE_BYTES, _stream())
    return g_tiles


# ---- GraspReadout on tcgen05 (delta_ngf/layers.py:8-42): forward + backward to the four stage outputs
def grasp_head_pack(params_flat: torch.Tensor) -> torch.Tensor:
    """Keras-order flat fp32 variables [HEAD_PARAMS] -> packed bf16 operand images + bias table"""
    params_flat = _f32c(params_flat)
    assert params_flat.numel() == _abi.HEAD_PARAMS
    pack = _bytes(_abi.HEAD_PACK_BYTES, params_flat.device)
    call('cn_grasp_head_pack', _ptr(params_flat), _ptr(pack), _stream())
    return pack


class HeadState:
    """what cn_grasp_head_fwd keeps for cn_grasp_head_bwd"""
    __slots__ = ('n_tiles', 'b', 'n_pad', 'np', 'pack', 'elu', 'vslabs', 'acat', 'saved')


def grasp_head_fwd(taps: torch.Tensor, b: int, n_pad: int, n_poses_per_scene: int, pack: torch.Tensor, elu: bool,
                   save: bool):
    """taps: bf16 activation-tile images uint8 [4, b * n_pad / 128 * ACT_TILE_BYTES] (``mlp_fusion_fwd_tap_tiles``),
    or fp32 rows [4, b * n_pad, 128] (converted with ``tile_add_rows``); rows '(b (np n5))' padded per scene to
    n_pad -> y [b, np], HeadState"""
    if taps.dtype == torch.float32:
        assert tuple(taps.shape) == (4, b * n_pad, 128)
        taps = torch.stack([tile_add_rows(None, taps[s]) for s in range(4)])
    assert taps.dtype == torch.uint8 and taps.is_contiguous() and tuple(taps.shape) == (4, b * n_pad // TILE_ROWS * ACT_TILE_BYTES)
    st = HeadState()
    st.n_tiles, st.b, st.n_pad, st.np, st.pack, st.elu = n_tiles(b * n_pad), b, n_pad, n_poses_per_scene, pack, int(elu)
    p_tiles = (b * n_poses_per_scene + 127) // 128
    dev = taps.device
    st.vslabs = _bytes(p_tiles * _abi.HEAD_OFFSETS * 2 * SLAB_BYTES, dev)
    st.acat = _bytes(st.n_tiles * 4 * SLAB_BYTES, dev) if save else None
    st.saved = _bytes(p_tiles * 5 * SLAB_BYTES, dev)
    y = torch.empty(b, n_poses_per_scene, dtype=torch.float32, device=dev)
    call('cn_grasp_head_fwd', _ptr(taps), st.n_tiles, b, n_pad, n_poses_per_scene, _ptr(pack), st.elu, _ptr(st.vslabs),
         _ptr(st.acat), _ptr(st.saved), _ptr(y), _stream())
    return y, st


def grasp_head_bwd(gy: torch.Tensor, st: HeadState) -> torch.Tensor:
    """gy [b, np] -> d y / d tap_s as bf16 activation-tile images, uint8 [4, n_tiles * ACT_TILE_BYTES]"""
    gy = _f32c(gy)
    assert gy.numel() == st.b * st.np and st.acat is not None
    p_tiles = (st.b * st.np + 127) // 128
    dpc = _bytes(p_tiles * _abi.HEAD_OFFSETS * SLAB_BYTES, gy.device)
    dtap = torch.empty(4, st.n_tiles * ACT_TILE_BYTES, dtype=torch.uint8, device=gy.device)
    call('cn_grasp_head_bwd', _ptr(gy), st.n_tiles, st.b, st.n_pad, st.np, _ptr(st.pack), st.elu, _ptr(st.vslabs),
         _ptr(st.acat), _ptr(st.saved), _ptr(dpc), _ptr(dtap), _stream())
    return dtap


def mlp_readout_bwd(params, d_y, acts, act_stride: int, act_index: int, grad):
    nt = d_y.numel() // (4 * TILE_ROWS)
    g = _bytes(nt * ACT_TILE_BYTES, d_y.device)
    call('cn_mlp_readout_bwd', _ptr(params), _ptr(d_y), _ptr(acts), act_stride, act_index, nt, _ptr(g), _ptr(grad),
         _stream())
    return g


def mlp_block_bwd(packed, blk: int, acts, act_stride: int, act_index: int, gy_tiles, grad, gh_tiles=None,
                  views: int = 1, scale: float = 1.0, b: int = 1):
    """views > 1: ``gy_tiles`` is the fusion-stage gradient [b][tiles_per_view], shared by the ``views`` per-view
    tiles of every row tile (mean over views, layers.py:368-370); the result goes to a new per-view buffer."""
    nt = gy_tiles.numel() // ACT_TILE_BYTES
    tpv = 0
    if views > 1:
        assert nt % b == 0
        tpv = nt // b       # gy_tiles is [b][tpv]; per-view tile (b, v, t) reads fusion tile (b, t)
        nt = nt * views
        assert gh_tiles is None
        gh_tiles = _bytes(nt * ACT_TILE_BYTES, gy_tiles.device)
    elif gh_tiles is None:
        gh_tiles = gy_tiles
    call('cn_mlp_block_bwd', _ptr(packed), blk, _ptr(acts), act_stride, act_index, nt, _ptr(gy_tiles), views, tpv,
         _ptr(gh_tiles), float(scale), _ptr(grad), _stream())
    return gh_tiles


def stage_bwd_workspace(device) -> torch.Tensor:
    """scratch of cn_mlp_stage_bwd (tile rings + counters, ~19 MiB: L2-resident)"""
    return _bytes(int(_abi.load().cn_mlp_stage_bwd_ws_bytes()), device)


def mlp_stage_bwd(packed, stage: int, acts, act_stride: int, gy_tiles, grad, ws=None, views: int = 1, b: int = 1,
                  scale: float = 1.0, x_tiles=None, want_dx: bool = False, first_dense: bool = False):
    """backward of a whole MLP stage in one launch (stage 1: fusion blocks; stage 0: per-view blocks and, with
    ``first_dense``, the first Dense's wgrad (+ dgrad when ``want_dx``)).  Returns the gradient w.r.t. the stage
    input (tile images; in place over ``gy_tiles`` when ``views == 1``) or, with ``first_dense``, dX (or None)."""
    nt = gy_tiles.numel() // ACT_TILE_BYTES
    tpv = 0
    if views > 1:
        assert nt % b == 0
        tpv = nt // b
        nt = nt * views
    if ws is None:
        ws = stage_bwd_workspace(gy_tiles.device)
    gh = dx = None
    if first_dense:
        assert stage == 0 and x_tiles is not None
        if want_dx:
            dx = _bytes(nt * X_TILE_BYTES, gy_tiles.device)
    else:
        gh = gy_tiles if views == 1 else _bytes(nt * ACT_TILE_BYTES, gy_tiles.device)
    call('cn_mlp_stage_bwd', _ptr(packed), stage, _ptr(acts), act_stride, nt, _ptr(gy_tiles), views, tpv, _ptr(gh),
         _ptr(x_tiles) if first_dense else None, _ptr(dx), float(scale), _ptr(grad), _ptr(ws), ws.numel(), _stream())
    return dx if first_dense else gh


def mlp_mean_bwd(gm_tiles, b: int, nv: int):
    nt = gm_tiles.numel() // ACT_TILE_BYTES
    tpv = nt // b
    out = _bytes(nt * nv * ACT_TILE_BYTES, gm_tiles.device)
    call('cn_mlp_mean_bwd', _ptr(gm_tiles), b, nv, tpv, _ptr(out), _stream())
    return out


def mlp_l0_dgrad(packed, g0_tiles, scale: float = 1.0):
    nt = g0_tiles.numel() // ACT_TILE_BYTES
    dx = _bytes(nt * X_TILE_BYTES, g0_tiles.device)
    call('cn_mlp_l0_dgrad', _ptr(packed), _ptr(g0_tiles), nt, _ptr(dx), float(scale), _stream())
    return dx


def mlp_l0_wgrad(x_tiles, g0_tiles, grad, scale: float = 1.0):
    nt = g0_tiles.numel() // ACT_TILE_BYTES
    call('cn_mlp_l0_wgrad', _ptr(x_tiles), _ptr(g0_tiles), nt, float(scale), _ptr(grad), _stream())


def adam_step(params, grad, m, v, n_update: int, step_counter, target_lr: float, warmup_steps: float,
              scale_down_after: float, grad_scale: float = 1.0, clip: float = 1.0, beta1: float = 0.9,
              beta2: float = 0.999, eps: float = 1e-7) -> None:
    """clip + warm-up schedule + Keras Adam on flat [n_rows, row_stride] fp32 buffers (first ``n_update`` of every row);
    ``step_counter``: 0-dim float64 CUDA tensor, incremented on the device"""
    for t in (params, grad, m, v):
        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.shape == params.shape
    assert step_counter.is_cuda and step_counter.dtype == torch.float64 and step_counter.numel() == 1
    n_rows, stride = (params.shape[0], params.shape[1]) if params.dim() == 2 else (1, params.numel())
    call('cn_adam_step', _ptr(params), _ptr(grad), _ptr(m), _ptr(v), n_rows, stride, int(n_update), _ptr(step_counter),
         float(target_lr), float(warmup_steps), float(scale_down_after), float(grad_scale), float(clip), float(beta1),
         float(beta2), float(eps), _stream())


def probe_tile_reduce(map_: torch.Tensor, iters: int, mode: int, ctas_per_sm: int = 1) -> None:
    n_img, H, W, C = map_.shape
    assert C == 256 and map_.dtype == torch.float32
    call('cn_probe_tile_reduce', _ptr(map_), n_img, H, W, iters, mode, ctas_per_sm, _stream())


def probe_umma(a_img: torch.Tensor, b_img: torch.Tensor, n_mma: int, a_mn: bool, b_mn: bool, a_lbo: int, a_sbo: int,
               a_kstep: int, b_lbo: int, b_sbo: int, b_kstep: int) -> torch.Tensor:
    d = torch.zeros(128, 128, dtype=torch.float32, device=a_img.device)
    call('cn_probe_umma', _ptr(a_img), a_img.numel(), _ptr(b_img), b_img.numel(), n_mma, int(a_mn), int(b_mn), a_lbo,
         a_sbo, a_kstep, b_lbo, b_sbo, b_kstep, _ptr(d), _stream())
    return d
