"""Forward-mode (tangent) pass of the pose-query embedding: d(taps) for a perturbation of the query
points / directions, with the ReLU masks of the primal pass held fixed.

Only ``DeltaNGF.train_step`` needs it (reference ``src/lib/delta_ngf/model.py:264-296``): the loss compares
d prediction / d pose with labelled directions and is differentiated w.r.t. the ``GraspReadout`` weights, so
autograd asks the backward of the query call for its own derivative w.r.t. the incoming tap gradients --
which is this Jacobian-vector product.  It runs on the same tcgen05 kernels as the forward:

  ``cn_query_tangent_x``        tangent of the MLP-input rows (projection, bilinear coordinates, posenc) as X tiles
  ``cn_mlp_feature_fwd_ex``     per-view stage on the tangent rows: every ReLU multiplies by the sign pattern of the
  ``cn_mlp_fusion_fwd_taps_ex`` activation tile the primal forward saved (``save_hidden``), biases dropped

``query_tangent_reference`` keeps the same computation written with torch ops on the kernels' saved tiles (library
GEMMs, index gathers): the GPU tests compare the two."""
from __future__ import annotations

from typing import List

import torch

from . import layout, ops
from . import params as P
from ._abi import TILE_ROWS

_PI32 = 3.14159274101257324


def query_tangent(scene: ops.Scene, packed: torch.Tensor, pts, drs, acts_f, acts_u, u_pts, u_dirs) -> List[torch.Tensor]:
    """J . (u_pts, u_dirs): pts, drs, u_* [b, n_pad, 3]; acts_f / acts_u: the activation tiles of a ``save_hidden``
    forward of the same rows -> 4 tangent taps [b, n_pad, 128]."""
    with torch.no_grad():
        b, n = pts.shape[:2]
        assert n % TILE_ROWS == 0
        x_dot = ops.query_tangent_x(scene, pts, drs, u_pts, u_dirs)
        from . import render as _render
        if scene.nv == 1 and _render.CHAIN_SINGLE_VIEW:
            _, _, _, taps = ops.mlp_chain_fwd(packed, x_dot, False, taps='f32', mask_acts_f=acts_f, mask_acts_u=acts_u)
        else:
            h3_dot, _ = ops.mlp_feature_fwd(packed, x_dot, False, mask_acts=acts_f)
            taps, _, _ = ops.mlp_fusion_fwd_taps(packed, h3_dot, b, scene.nv, False, mask_acts=acts_u)
        return list(taps.view(4, b, n, 128).unbind(0))


def _posenc_tangent(v: torch.Tensor, v_dot: torch.Tensor) -> torch.Tensor:
    """d/dt position_encoding(v + t v_dot) (nerf_utils.py:108-126): [..., 3] -> [..., 60] in (c, k, sin|cos) order"""
    f = (_PI32 * torch.pow(2.0, torch.arange(10, device=v.device, dtype=torch.float32))).to(torch.float32)
    arg = v[..., None] * f
    s = f * v_dot[..., None]
    return torch.stack([torch.cos(arg) * s, -torch.sin(arg) * s], dim=-1).reshape(*v.shape[:-1], 60)


def _q(w: torch.Tensor) -> torch.Tensor:
    """the operand the tensor cores see: bf16-rounded, held in fp32"""
    return w.to(torch.bfloat16).to(torch.float32)


def _block_tangent(h_dot, relu_h, relu_t, wa, wb):
    """tangent of y = h + relu(relu(h) Wa + ba) Wb + bb with the primal masks (layers.py:262-298)"""
    return h_dot + (((h_dot * (relu_h > 0).to(h_dot.dtype)) @ wa) * (relu_t > 0).to(h_dot.dtype)) @ wb


def query_tangent_reference(scene: ops.Scene, params: torch.Tensor, pts, drs, z, acts_f, acts_u, u_pts, u_dirs
                            ) -> List[torch.Tensor]:
    """the same Jacobian-vector product with torch ops (fp32 GEMMs on bf16-rounded weights) -- test reference"""
    with torch.no_grad():
        b, nv, H, W = scene.b, scene.nv, scene.H, scene.W
        n = pts.shape[1]
        E, K = scene.extr_inv, scene.intrinsics
        pix, corner, _, cam = ops.project_gather_debug(scene, pts, drs, z, gather=False, cam=True)
        pix, corner, cam = pix[:, :, :, 0], corner[:, :, :, 0].long(), cam[:, :, :, 0]        # [b, nv, n, .]
        E3 = E[:, :, :3, :3]
        p_dot = torch.einsum('bvij,bnj->bvni', E3, u_pts)
        d_dot = torch.einsum('bvij,bnj->bvni', E3, u_dirs)
        dcam = torch.einsum('bvij,bnj->bvni', E3, drs) + E[:, :, None, :3, 3]                # E [d; 1] (quirk Q1)
        q = torch.einsum('bvij,bvnj->bvni', K[:, :, :3, :], cam)
        q_dot = torch.einsum('bvij,bvnj->bvni', K[:, :, :3, :3], p_dot)
        den = torch.clamp(q[..., 2], min=1e-8)
        rx, ry = q[..., 0] / den, q[..., 1] / den
        qz_dot = torch.where(q[..., 2] >= 1e-8, q_dot[..., 2], torch.zeros_like(den))
        tx, ty = pix[..., 0] - corner[..., 0], pix[..., 1] - corner[..., 1]
        ok_x = (tx >= 0) & (tx <= 1) & (rx >= -1e6) & (rx <= 1e6)
        ok_y = (ty >= 0) & (ty <= 1) & (ry >= -1e6) & (ry <= 1e6)
        px_dot = torch.where(ok_x, (q_dot[..., 0] - rx * qz_dot) / den, torch.zeros_like(den))
        py_dot = torch.where(ok_y, (q_dot[..., 1] - ry * qz_dot) / den, torch.zeros_like(den))
        ax, ay = tx.clamp(0, 1)[..., None], ty.clamp(0, 1)[..., None]
        bv = torch.arange(b * nv, device=pts.device).view(b, nv, 1)
        i00 = ((bv * H + corner[..., 1]) * W + corner[..., 0]).reshape(-1)

        def lerp_tangent(grid, c, scale):
            g = grid.reshape(-1, c)
            tl, tr, bl, br = g[i00], g[i00 + 1], g[i00 + W], g[i00 + W + 1]
            axf, ayf = ax.reshape(-1, 1), ay.reshape(-1, 1)
            gx = (1 - ayf) * (tr - tl) + ayf * (br - bl)
            gy = (1 - axf) * (bl - tl) + axf * (br - tr)
            return scale * (px_dot.reshape(-1, 1) * gx + py_dot.reshape(-1, 1) * gy)

        x_dot = torch.cat([_posenc_tangent(cam[..., :3], p_dot).reshape(-1, 60),
                           _posenc_tangent(dcam, d_dot).reshape(-1, 60), lerp_tangent(scene.images, 3, 2.0),
                           lerp_tangent(scene.features, 256, 1.0)], dim=-1)                       # layers.py:362 order
        w = P.views(params)
        h_dot = x_dot @ _q(w['embedding.layer_0.kernel'])
        rows_v = b * nv * n
        a_f = layout.tiles_to_matrix(acts_f, 128).view(rows_v // TILE_ROWS, ops.FEATURE_ACTS + 3, TILE_ROWS, 128)
        for i in range(3):
            pre = f'embedding.feature_blocks.{i}'
            h_dot = _block_tangent(h_dot, a_f[:, i].reshape(rows_v, 128), a_f[:, 3 + i].reshape(rows_v, 128),
                                   _q(w[f'{pre}.layer_0.kernel']), _q(w[f'{pre}.layer_1.kernel']))
        taps = [h_dot.view(b, nv, n, 128).mean(dim=1).reshape(b * n, 128)]
        a_u = layout.tiles_to_matrix(acts_u, 128).view(b * n // TILE_ROWS, ops.FUSION_ACTS + 3, TILE_ROWS, 128)
        for i in range(3):
            pre = f'embedding.fusion_blocks.{i}'
            taps.append(_block_tangent(taps[-1], a_u[:, i].reshape(b * n, 128), a_u[:, 4 + i].reshape(b * n, 128),
                                       _q(w[f'{pre}.layer_0.kernel']), _q(w[f'{pre}.layer_1.kernel'])))
        return [t.view(b, n, 128) for t in taps]
