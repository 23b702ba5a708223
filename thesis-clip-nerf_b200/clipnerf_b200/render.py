"""Differentiable CLIP-NeRF render call on the sm_100a kernels.

``render_rays`` is the replacement of ``MVVNeRFRenderer._call`` (reference
``src/lib/mvnerf/model_v0.py:113-184``): coarse pass (stratified samples -> project / gather /
posenc -> per-view MLP -> view mean + fusion MLP + readout -> composite), hierarchical resampling,
fine pass; the backward walks the same chain in reverse, including the gradient that flows from the
fine samples through the inverse-CDF sampler into the coarse net (no stop_gradient, quirk Q4).

Random draws are explicit inputs (``u_strat``, ``u_pdf``)."""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from . import ops
from ._abi import NET_PARAMS, TILE_ROWS


class _Pass:
    """Everything one (coarse or fine) pass keeps for its backward."""
    __slots__ = ('z', 'x', 'ws', 'acts_f', 'acts_u', 'rgbs', 'rgb', 'depth', 'weights', 'packed')


def _pass_fwd(packed, scene: ops.Scene, ray_o, ray_d, z, save: bool) -> _Pass:
    P = _Pass()
    P.z, P.packed = z, packed
    P.x, P.ws = ops.project_gather_fwd(scene, ray_o, ray_d, z, return_ws=True)
    if scene.nv == 1 and CHAIN_SINGLE_VIEW:
        # one source view: both MLP stages + readout in one launch (cn_mlp_chain_fwd), same bits
        P.rgbs, P.acts_f, P.acts_u, _ = ops.mlp_chain_fwd(packed, P.x, save)
    else:
        h3, P.acts_f = ops.mlp_feature_fwd(packed, P.x, save)
        P.rgbs, P.acts_u = ops.mlp_fusion_fwd(packed, h3, scene.b, scene.nv, save)
    P.rgb, P.depth, P.weights = ops.composite_fwd(z, P.rgbs.view(*z.shape, 4))
    if not save:
        P.x = P.ws = None
    return P


def _pass_bwd(P: _Pass, params, grad, scene: ops.Scene, ray_o, ray_d, d_rgb, d_depth, d_weights, want_dx: bool,
              need_dz: bool):
    """backward of one pass up to dX (the gradient of the MLP input rows); the feature-map gradient of both
    passes is taken from the two dX images in one sweep afterwards (ops.feature_grad)."""
    d_y, dz = ops.composite_bwd(P.z, P.rgbs.view(*P.z.shape, 4), d_rgb, d_depth, d_weights, need_dz)
    g = ops.mlp_readout_bwd(params, d_y, P.acts_u, ops.FUSION_ACTS, 3, grad)
    if CHAINED_STAGES:
        # one launch per stage: role-specialised CTAs pass the gradient stream through L2-resident tile rings
        ws = ops.stage_bwd_workspace(g.device)
        ops.mlp_stage_bwd(P.packed, 1, P.acts_u, ops.FUSION_ACTS, g, grad, ws)
        dx = ops.mlp_stage_bwd(P.packed, 0, P.acts_f, ops.FEATURE_ACTS, g, grad, ws, views=scene.nv, b=scene.b,
                               scale=1.0 / scene.nv, x_tiles=P.x, want_dx=want_dx or need_dz, first_dense=True)
        if need_dz:
            ops.project_gather_bwd(scene, ray_o, ray_d, P.z, dx, None, dz, P.ws)
        return dz, dx
    for blk in (5, 4, 3):
        ops.mlp_block_bwd(P.packed, blk, P.acts_u, ops.FUSION_ACTS, blk - 3, g, grad)
    # mean over views: every per-view tile reads the fusion-stage gradient tile itself; the factor 1/nv is
    # applied in fp32 wherever the per-view stage leaves the bf16 g stream (weight / bias gradients, dX)
    s = 1.0 / scene.nv
    gf = ops.mlp_block_bwd(P.packed, 2, P.acts_f, ops.FEATURE_ACTS, 2, g, grad, views=scene.nv, scale=s, b=scene.b) \
        if scene.nv > 1 else ops.mlp_block_bwd(P.packed, 2, P.acts_f, ops.FEATURE_ACTS, 2, g, grad)
    for blk in (1, 0):
        ops.mlp_block_bwd(P.packed, blk, P.acts_f, ops.FEATURE_ACTS, blk, gf, grad, scale=s)
    ops.mlp_l0_wgrad(P.x, gf, grad, s)
    dx = None
    if want_dx or need_dz:
        dx = ops.mlp_l0_dgrad(P.packed, gf, s)
        if need_dz:
            ops.project_gather_bwd(scene, ray_o, ray_d, P.z, dx, None, dz, P.ws)
    return dz, dx


# True: the MLP backward of a pass is two launches (cn_mlp_stage_bwd per stage, role-specialised CTAs chained through
# L2 tile rings: 45 % less DRAM traffic); False: one launch per block / Dense, every SM on every tile.  Measured on the
# cfg2 step the two are within 2 % (profiles/README.md, round 2), the per-block launches ahead: the block backward is
# bound by shared-memory bandwidth inside the SM, not by HBM, so the default stays False.
CHAINED_STAGES = os.environ.get('CN_CHAINED_STAGES', '0') == '1'

# nv = 1: chain the fusion stage onto the per-view stage in one kernel (False: the two-kernel path, for comparison)
CHAIN_SINGLE_VIEW = os.environ.get('CN_CHAIN_SINGLE_VIEW', '1') == '1'

# Set by RenderTrainer for the duration of a backward on > 1 rank: callable([2, NET_PARAMS] gradient buffer) -> a
# work handle with .wait() (torch.distributed all_reduce(async_op=True)); None = no collective inside the backward.
GRAD_SYNC = None


class _RenderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, params_c, params_f, features, images, intrinsics, extr_inv, ray_o, ray_d, u_strat, u_pdf,
                near, far):
        need_grad = any(ctx.needs_input_grad[:3])
        scene = ops.Scene(images, features, intrinsics, extr_inv)
        pk_c = ops.pack_weights(params_c)
        pk_f = ops.pack_weights(params_f)
        z_c = ops.sample_coarse(u_strat, near, far)
        C = _pass_fwd(pk_c, scene, ray_o, ray_d, z_c, need_grad)
        z_f, perm, _above = ops.sample_pdf_fwd(z_c, C.weights, u_pdf)
        Fp = _pass_fwd(pk_f, scene, ray_o, ray_d, z_f, need_grad)
        if need_grad:
            ctx.scene, ctx.C, ctx.F = scene, C, Fp
            ctx.perm, ctx.u_pdf = perm, ops._f32c(u_pdf)
            ctx.rays = (ops._f32c(ray_o), ops._f32c(ray_d))
            ctx.save_for_backward(params_c, params_f)
        ctx.mark_non_differentiable()
        return C.rgb, C.depth, Fp.rgb, Fp.depth

    @staticmethod
    def backward(ctx, d_rgb_c, d_depth_c, d_rgb_f, d_depth_f):
        params_c, params_f = ctx.saved_tensors
        scene, C, Fp = ctx.scene, ctx.C, ctx.F
        ray_o, ray_d = ctx.rays
        dev = params_c.device
        need_feat = ctx.needs_input_grad[2]
        g_both = torch.zeros(2, NET_PARAMS, dtype=torch.float32, device=dev)   # one buffer: one collective (GRAD_SYNC)
        g_c, g_f = g_both[0], g_both[1]
        # written exactly once per texel by ops.feature_grad: no zero fill
        d_features = torch.empty_like(scene.features) if need_feat else None

        def z_or_none(t):
            return None if t is None else t.contiguous()

        d_rgb_f = torch.zeros_like(Fp.rgb) if d_rgb_f is None else d_rgb_f
        d_rgb_c = torch.zeros_like(C.rgb) if d_rgb_c is None else d_rgb_c
        # fine pass first: its dz feeds the coarse weights through the inverse-CDF sampler
        dz_f, dx_f = _pass_bwd(Fp, params_f, g_f, scene, ray_o, ray_d, d_rgb_f, z_or_none(d_depth_f), None, need_feat,
                               True)
        dw_c = ops.sample_pdf_bwd(C.z, C.weights, ctx.u_pdf, ctx.perm, dz_f)
        _, dx_c = _pass_bwd(C, params_c, g_c, scene, ray_o, ray_d, d_rgb_c, z_or_none(d_depth_c), dw_c, need_feat, False)
        # both nets' parameter gradients are complete here; the feature-map sweep below (the longest kernel of the
        # step) does not touch them, so the data-parallel all-reduce is started now and runs under it
        work = GRAD_SYNC(g_both) if GRAD_SYNC is not None else None
        if need_feat:
            nr = C.z.shape[1]
            ops.feature_grad(scene, [(Fp.ws, dx_f, nr, Fp.z.shape[2]), (C.ws, dx_c, nr, C.z.shape[2])], d_features,
                             accumulate=False)
        if work is not None:
            work.wait()                                   # the compute stream waits, the host does not
        ctx.C = ctx.F = ctx.scene = None
        return (g_c, g_f, d_features) + (None,) * 9


def render_rays(params_coarse: torch.Tensor, params_fine: torch.Tensor, inputs, combined_features: torch.Tensor,
                u_strat: torch.Tensor, u_pdf: torch.Tensor, near: float, far: float
                ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    """(rgb [b,nr,3], depth [b,nr], fine_rgb, fine_depth) = _call(inputs, features).

    inputs: the 5-tuple of ``MVNeRFDataGenerator.get_input`` as CUDA tensors
    (ray_o [b,nr,3], ray_d [b,nr,3], images [b,nv,H,W,3] in [0,1], K [b,nv,4,4], E^-1 [b,nv,4,4]);
    ``params_*``: flat parameter vectors (``params.py``); n_samples is fixed at 64 coarse + 128 fine
    (``u_strat`` / ``u_pdf`` are [b, nr, 64]).  Any nr: the kernels work on tiles of 128 rows = 2 rays of 64 coarse
    samples, an odd batch is rendered with its last ray repeated (the copy's outputs are dropped and receive no
    gradient); an empty batch returns empty outputs without a launch."""
    ray_o, ray_d, images, K, Einv = inputs
    assert u_strat.shape[-1] == 64 and u_pdf.shape[-1] == 64, 'n_samples is fixed at 64 (sample_pdf kernel)'
    b, nr = ray_o.shape[:2]
    if nr == 0:
        return ray_o.new_zeros(b, 0, 3), ray_o.new_zeros(b, 0), ray_o.new_zeros(b, 0, 3), ray_o.new_zeros(b, 0)
    if nr % 2:
        ray_o, ray_d, u_strat, u_pdf = [torch.cat([t, t[:, -1:]], dim=1) for t in (ray_o, ray_d, u_strat, u_pdf)]
    out = _RenderFn.apply(params_coarse, params_fine, combined_features, images, K, Einv, ray_o, ray_d, u_strat,
                          u_pdf, float(near), float(far))
    return tuple(o[:, :nr] for o in out) if nr % 2 else out
