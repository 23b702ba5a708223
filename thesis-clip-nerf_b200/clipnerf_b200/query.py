"""Differentiable pose-query call on the sm_100a kernels.

``query_embeddings`` replaces everything of ``DeltaNGF._call`` / ``LanguageNeRF._call`` /
``GraspMVNeRF._call`` between the pose matrices and the ``GraspReadout`` (reference
``src/lib/delta_ngf/model.py:188-241``, ``lmvnerf/model_v4.py:208-261``,
``grasp_mvnerf/model.py:113-155``): the 42 fixed query offsets of every pose are projected into the
source views, the 259-channel map is gathered bilinearly, points and approach directions are
position-encoded and the fine embedding runs with ``complete_output`` -- the same gather and
tcgen05 MLP kernels as the render path, on rows with one sample per "ray" (S = 1, z = 0,
origin = query point, direction = approach axis).  The backward returns the gradient with respect
to the query points and directions (what ``DNGFOptimizer.optimize_pose`` and the gradient-matching
loss of ``DeltaNGF.train_step`` differentiate), the embedding parameters and the feature maps."""
from __future__ import annotations

from typing import List

import torch

from . import ops
from . import render as _render
from ._abi import HEAD_OFFSETS as _HEAD_OFFSETS, NET_PARAMS, TILE_ROWS


def _pad_rows(t: torch.Tensor, n_pad: int) -> torch.Tensor:
    """[b, n, c] -> [b, n_pad, c]; the padding repeats row 0 (any valid point: its outputs are dropped)"""
    b, n, c = t.shape
    if n == n_pad:
        return t.contiguous()
    return torch.cat([t, t[:, :1].expand(b, n_pad - n, c)], dim=1).contiguous()


class _State:
    """what the forward keeps for the backward and for the tangent pass of the double backward"""
    __slots__ = ('scene', 'params', 'packed', 'n', 'pts', 'drs', 'z', 'x', 'ws', 'acts_f', 'acts_u', 'need', 'fa', 'ua')


class _QueryEmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, params, features, images, intrinsics, extr_inv, points, dirs):
        need_grad = any(ctx.needs_input_grad)
        scene = ops.Scene(images, features, intrinsics, extr_inv)
        b, n, _ = points.shape
        n_pad = (n + TILE_ROWS - 1) // TILE_ROWS * TILE_ROWS
        pts, drs = _pad_rows(ops._f32c(points), n_pad), _pad_rows(ops._f32c(dirs), n_pad)
        z = torch.zeros(b, n_pad, 1, dtype=torch.float32, device=pts.device)
        packed = ops.pack_weights(params)
        x, ws = ops.project_gather_fwd(scene, pts, drs, z, return_ws=True)
        # the hidden activations are kept as well (3 more tiles per tile): the tangent pass of a double backward
        # (DeltaNGF.train_step) takes its ReLU patterns from them
        if scene.nv == 1 and _render.CHAIN_SINGLE_VIEW:
            _rgbs, acts_f, acts_u, taps = ops.mlp_chain_fwd(packed, x, need_grad, taps='f32', save_hidden=need_grad)
        else:
            h3, acts_f = ops.mlp_feature_fwd(packed, x, need_grad, save_hidden=need_grad)
            taps, _rgbs, acts_u = ops.mlp_fusion_fwd_taps(packed, h3, b, scene.nv, need_grad, save_hidden=need_grad)
        if need_grad:
            st = _State()
            st.fa, st.ua = ops.FEATURE_ACTS + 3, ops.FUSION_ACTS + 3      # act_stride_tiles of the backward kernels
            st.scene, st.params, st.packed, st.n = scene, params.detach(), packed, n
            st.pts, st.drs, st.z, st.x, st.ws, st.acts_f, st.acts_u = pts, drs, z, x, ws, acts_f, acts_u
            st.need = tuple(ctx.needs_input_grad)
            ctx.state = st
        return tuple(taps.view(4, b, n_pad, 128)[:, :, :n].unbind(0))

    @staticmethod
    def backward(ctx, *g_taps):
        st = ctx.state
        b, n_pad = st.z.shape[:2]
        g_taps = [torch.zeros(b, st.n, 128, dtype=torch.float32, device=st.z.device) if g is None else g for g in g_taps]
        # a Function of its own, so that autograd can differentiate THIS backward w.r.t. the incoming tap
        # gradients (create_graph=True in DeltaNGF.train_step)
        d_pts, d_dirs, grad, d_features = _QueryBwdFn.apply(st, *g_taps)
        need = st.need
        return (grad if need[0] else None), (d_features if need[1] else None), None, None, None, \
            (d_pts if need[5] else None), (d_dirs if need[6] else None)


class _QueryBwdFn(torch.autograd.Function):
    """(g_tap0..3) -> (d_points, d_dirs, d_params, d_features) on the CUDA kernels; its own backward is the
    Jacobian-vector product of the forward (``query_tangent``)."""

    @staticmethod
    def forward(ctx, st, *g_taps):
        scene, packed, n = st.scene, st.packed, st.n
        b, n_pad = st.z.shape[:2]
        dev = st.z.device
        need = st.need
        grad = torch.zeros(NET_PARAMS, dtype=torch.float32, device=dev)

        def rows(i):
            g = g_taps[i].detach().to(torch.float32)
            if n_pad != n:
                g = torch.cat([g, torch.zeros(b, n_pad - n, 128, dtype=torch.float32, device=dev)], dim=1)
            return g.reshape(b * n_pad, 128).contiguous()

        # the gradient of every tapped stage output joins the bf16 gradient stream where the stage was produced
        g = ops.tile_add_rows(None, rows(3))
        for blk in (5, 4, 3):
            ops.mlp_block_bwd(packed, blk, st.acts_u, st.ua, blk - 3, g, grad)
            ops.tile_add_rows(g, rows(blk - 3))
        s = 1.0 / scene.nv
        gf = ops.mlp_block_bwd(packed, 2, st.acts_f, st.fa, 2, g, grad, views=scene.nv, scale=s, b=scene.b) \
            if scene.nv > 1 else ops.mlp_block_bwd(packed, 2, st.acts_f, st.fa, 2, g, grad)
        for blk in (1, 0):
            ops.mlp_block_bwd(packed, blk, st.acts_f, st.fa, blk, gf, grad, scale=s)
        if need[0]:
            ops.mlp_l0_wgrad(st.x, gf, grad, s)
        d_features = torch.empty(0, device=dev)
        d_pts = torch.zeros(b, n, 3, dtype=torch.float32, device=dev)
        d_dirs = torch.zeros(b, n, 3, dtype=torch.float32, device=dev)
        if need[1] or need[5] or need[6]:
            dx = ops.mlp_l0_dgrad(packed, gf, s)
            if need[5] or need[6]:
                d_o, d_d, _ = ops.project_gather_bwd_rays(scene, st.pts, st.drs, st.z, dx, st.ws)
                d_pts, d_dirs = d_o[:, :n].contiguous(), d_d[:, :n].contiguous()
            if need[1]:
                d_features = torch.empty_like(scene.features)
                ops.feature_grad(scene, [(st.ws, dx, n_pad, 1)], d_features, accumulate=False)
        ctx.state = st
        ctx.mark_non_differentiable(grad, d_features)
        return d_pts, d_dirs, grad, d_features

    @staticmethod
    def backward(ctx, u_pts, u_dirs, _ug, _uf):
        from .query_tangent import query_tangent
        st = ctx.state
        b, n_pad = st.z.shape[:2]
        n = st.n

        def pad(u):
            u = torch.zeros(b, n, 3, dtype=torch.float32, device=st.z.device) if u is None else u.to(torch.float32)
            return torch.cat([u, u.new_zeros(b, n_pad - n, 3)], dim=1) if n_pad != n else u
        taps = query_tangent(st.scene, st.packed, st.pts, st.drs, st.acts_f, st.acts_u, pad(u_pts), pad(u_dirs))
        return (None,) + tuple(t[:, :n] for t in taps)


class _QueryGraspFn(torch.autograd.Function):
    """query points -> grasp success with the affordance head on the CUDA kernels too (``cn_grasp_head_fwd`` /
    ``_bwd``): the path of ``infer`` and of ``DNGFOptimizer.optimize_pose`` (grasp_optimizer.py:158-184), where the
    field and the head are frozen and only d success / d (points, directions) is wanted.  The head's tap gradients
    come back as bf16 tiles and join the gradient stream of the block backward directly.  Once differentiable;
    training the head (``DeltaNGF.train_step``, double backward) uses ``query_embeddings`` + the torch head."""

    @staticmethod
    def forward(ctx, params, head_pack, features, images, intrinsics, extr_inv, points, dirs, n_poses, elu):
        need_grad = ctx.needs_input_grad[6] or ctx.needs_input_grad[7]
        scene = ops.Scene(images, features, intrinsics, extr_inv)
        b, n, _ = points.shape
        n_pad = (n + TILE_ROWS - 1) // TILE_ROWS * TILE_ROWS
        pts, drs = _pad_rows(ops._f32c(points), n_pad), _pad_rows(ops._f32c(dirs), n_pad)
        z = torch.zeros(b, n_pad, 1, dtype=torch.float32, device=pts.device)
        packed = ops.pack_weights(params)
        x, ws = ops.project_gather_fwd(scene, pts, drs, z, return_ws=True)
        if scene.nv == 1 and _render.CHAIN_SINGLE_VIEW:
            _rgbs, acts_f, acts_u, taps = ops.mlp_chain_fwd(packed, x, need_grad, taps='bf16')
        else:
            h3, acts_f = ops.mlp_feature_fwd(packed, x, need_grad)
            taps, _rgbs, acts_u = ops.mlp_fusion_fwd_tap_tiles(packed, h3, b, scene.nv, need_grad)
        y, head = ops.grasp_head_fwd(taps, b, n_pad, n_poses, head_pack, elu, need_grad)
        if need_grad:
            st = _State()
            st.fa, st.ua = ops.FEATURE_ACTS, ops.FUSION_ACTS
            st.scene, st.params, st.packed, st.n = scene, None, packed, n
            st.pts, st.drs, st.z, st.x, st.ws, st.acts_f, st.acts_u = pts, drs, z, None, ws, acts_f, acts_u
            st.need = tuple(ctx.needs_input_grad)
            ctx.state, ctx.head = st, head
        return y

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gy):
        st, head = ctx.state, ctx.head
        scene, packed, n = st.scene, st.packed, st.n
        dev = st.z.device
        dtap = ops.grasp_head_bwd(gy.contiguous(), head)
        grad = None                                   # frozen field: the block backwards form no weight gradients
        g = dtap[3]
        for blk in (5, 4, 3):
            ops.mlp_block_bwd(packed, blk, st.acts_u, ops.FUSION_ACTS, blk - 3, g, grad)
            ops.tile_add_tiles(g, dtap[blk - 3])
        s = 1.0 / scene.nv
        gf = ops.mlp_block_bwd(packed, 2, st.acts_f, ops.FEATURE_ACTS, 2, g, grad, views=scene.nv, scale=s, b=scene.b) \
            if scene.nv > 1 else ops.mlp_block_bwd(packed, 2, st.acts_f, ops.FEATURE_ACTS, 2, g, grad)
        for blk in (1, 0):
            ops.mlp_block_bwd(packed, blk, st.acts_f, ops.FEATURE_ACTS, blk, gf, grad, scale=s)
        dx = ops.mlp_l0_dgrad(packed, gf, s)
        d_o, d_d, _ = ops.project_gather_bwd_rays(scene, st.pts, st.drs, st.z, dx, st.ws)
        need = st.need
        return (None,) * 6 + ((d_o[:, :n].contiguous() if need[6] else None),
                              (d_d[:, :n].contiguous() if need[7] else None), None, None)


def query_grasp_success(params_fine: torch.Tensor, head_pack: torch.Tensor, images, intrinsics, extr_inv,
                        combined_features, points, dirs, n_poses: int, elu: bool = True) -> torch.Tensor:
    """grasp success [b, n_poses] of the poses whose 42 query points / directions are ``points`` / ``dirs``
    [b, n_poses * 42, 3]: embedding + ``GraspReadout`` on the CUDA kernels, differentiable w.r.t. points and dirs."""
    assert points.shape == dirs.shape and points.shape[1] == n_poses * _HEAD_OFFSETS
    return _QueryGraspFn.apply(params_fine, head_pack, combined_features, images, intrinsics, extr_inv, points, dirs,
                               int(n_poses), bool(elu))


def query_embeddings(params_fine: torch.Tensor, images, intrinsics, extr_inv, combined_features, points, dirs
                     ) -> List[torch.Tensor]:
    """``fine_embedding([camera_points, camera_directions, gathered])[4:]`` for world-frame query points.

    params_fine: flat parameter vector of one net (``params.py``; the RenderReadout tail is unused);
    images [b,nv,H,W,3] in [0,1], intrinsics / extr_inv [b,nv,4,4], combined_features [b,nv,H,W,256];
    points, dirs [b, n, 3] (world frame).  Returns 4 tensors [b, n, 128] (mean over views and the outputs
    of the three fusion blocks)."""
    assert points.shape == dirs.shape and points.shape[-1] == 3
    return list(_QueryEmbedFn.apply(params_fine, combined_features, images, intrinsics, extr_inv, points, dirs))


class _PosePointsFn(torch.autograd.Function):
    """transforms [b, np, 4, 4] -> (points, dirs) [b, np * n5, 3] on the device kernels (cn_pose_points_fwd / _bwd);
    its backward is linear in the incoming gradients and written with the same kernel, so the double backward of
    DeltaNGF.train_step passes through it."""

    @staticmethod
    def forward(ctx, transforms, ttc):
        b, n_poses = transforms.shape[:2]
        n5 = ttc.shape[0]
        M, ttc = ops._f32c(transforms), ops._f32c(ttc)
        pts = torch.empty(b, n_poses * n5, 3, dtype=torch.float32, device=M.device)
        dirs = torch.empty_like(pts)
        ops.call('cn_pose_points_fwd', ops._ptr(M), ops._ptr(ttc), b * n_poses, n5, ops._ptr(pts), ops._ptr(dirs), ops._stream())
        ctx.save_for_backward(ttc)
        ctx.shape = (b, n_poses, n5)
        return pts, dirs

    @staticmethod
    def backward(ctx, d_pts, d_dirs):
        ttc, = ctx.saved_tensors
        return _PosePointsBwdFn.apply(d_pts, d_dirs, ttc, ctx.shape), None


class _PosePointsBwdFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, d_pts, d_dirs, ttc, shape):
        b, n_poses, n5 = shape
        z = None
        if d_pts is None or d_dirs is None:
            z = torch.zeros(b, n_poses * n5, 3, dtype=torch.float32, device=ttc.device)
        d_pts = ops._f32c(d_pts) if d_pts is not None else z
        d_dirs = ops._f32c(d_dirs) if d_dirs is not None else z
        dM = torch.empty(b, n_poses, 4, 4, dtype=torch.float32, device=ttc.device)
        ops.call('cn_pose_points_bwd', ops._ptr(d_pts), ops._ptr(d_dirs), ops._ptr(ttc), b * n_poses, n5, ops._ptr(dM),
                 ops._stream())
        ctx.save_for_backward(ttc)
        return dM

    @staticmethod
    def backward(ctx, u):
        # dM = sum_j d_pt c_j^T + d_dir z_j^T is linear in (d_pts, d_dirs): its adjoint is the forward map applied to u
        ttc, = ctx.saved_tensors
        u = ops._f32c(u)                                  # (the kernel reads rows 0..2 of every matrix only)
        b, n_poses = u.shape[:2]
        n5 = ttc.shape[0]
        g_pts = torch.empty(b, n_poses * n5, 3, dtype=torch.float32, device=u.device)
        g_dirs = torch.empty_like(g_pts)
        ops.call('cn_pose_points_fwd', ops._ptr(u), ops._ptr(ttc), b * n_poses, n5, ops._ptr(g_pts), ops._ptr(g_dirs),
                 ops._stream())
        return g_pts, g_dirs, None, None


def pose_query_points(transforms: torch.Tensor, transforms_to_check: torch.Tensor):
    """poses = transforms . transforms_to_check (delta_ngf/model.py:199-200, 221): transforms [b, np, 4, 4],
    transforms_to_check [n5, 4, 4] -> query points and approach directions [b, np * n5, 3] in the row order
    '(np n5)' the embedding / GraspReadout use (model.py:228-238): point = R c_j + t, direction = R (C_j[:3, :3] z)
    (the bottom row of M is [0 0 0 1], C_j[3, 3] = 1).  CUDA tensors go through cn_pose_points_fwd / _bwd."""
    if transforms.is_cuda:
        return _PosePointsFn.apply(transforms, transforms_to_check.to(transforms.device))
    ttc = transforms_to_check.to(transforms.dtype)
    R, t = transforms[:, :, :3, :3], transforms[:, :, :3, 3]
    pts = (R[:, :, None] * ttc[None, None, :, None, :3, 3]).sum(-1) + t[:, :, None]
    dirs = (R[:, :, None] * ttc[None, None, :, None, :3, 2]).sum(-1)
    b, n_pts, n5 = pts.shape[:3]
    return pts.reshape(b, n_pts * n5, 3), dirs.reshape(b, n_pts * n5, 3)
