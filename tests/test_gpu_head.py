"""Parity of the tcgen05 affordance head (cn_grasp_head_fwd / _bwd) with the oracle's ``grasp_readout``
(reference ``src/lib/delta_ngf/layers.py:8-42``; relu variant ``grasp_mvnerf/layers.py``): grasp success
(rel-err <= 1e-2) and the gradient w.r.t. the four tapped stage outputs (per-row statistics as for the pose
gradients, DESIGN.md section 2: p90 <= 2e-2, relative L2 <= 5e-2), plus the module-level frozen-field path of
``DeltaNGF._call`` / ``DNGFOptimizer`` that uses it."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _flat(named, O):
    parts = [named[f'grasp_readout.{n}'].reshape(-1) for n, _ in O.grasp_readout_param_shapes(True)]
    return torch.cat(parts).float()


def _row_stats(got, ref):
    got, ref = got.detach().double().cpu(), ref.detach().double().cpu()
    got, ref = got.reshape(-1, got.shape[-1]), ref.reshape(-1, ref.shape[-1])
    e = (got - ref).norm(dim=-1) / ref.norm(dim=-1).max().clamp_min(1e-30)
    return float(e.median()), float(torch.quantile(e, 0.9)), float(e.max()), float((got - ref).norm() / ref.norm())


def _unflipped_poses(st, o2, tag):
    """A pose whose final ReLU input has an element within the accumulation noise of zero (fp32 tensor-core sums vs
    the oracle's fp64) can come out with the other mask: ONE flipped element changes the gradient of all 42 rows of
    the pose by ~ |w_o[k]| / |w_o| ("ReLU masks that flip").  The kernel's own mask is read back
    from the tile it saves for the backward (relu(o2), slab 4): poses whose mask differs from the oracle's are
    excluded from the gradient statistics -- after checking that every flipped element really is a near-zero one."""
    from clipnerf_b200 import layout
    o2 = o2.detach().double().cpu()
    kern = layout.tiles_to_matrix(st.saved, 320)[:o2.shape[0], 256:320].cpu() > 0
    flipped = kern != (o2 > 0)
    assert float(o2.abs()[flipped].max() if flipped.any() else 0.0) < 1e-2 * float(o2.abs().max()), 'a mask differs far from zero'
    clear = ~flipped.any(dim=-1)
    print(f'[{tag}] poses with a flipped final-ReLU mask element: {int((~clear).sum())} of {clear.numel()}')
    assert int((~clear).sum()) <= max(1, clear.numel() // 4)
    return clear


@pytest.mark.parametrize('b,n_poses,act', [(1, 3, 'elu'), (2, 5, 'relu'), (2, 200, 'elu'), (3, 130, 'elu')])
def test_grasp_head_forward_and_tap_gradients(cuda, b, n_poses, act):
    from clipnerf_b200 import _abi, layout, ops
    from oracle import nerf_oracle as O
    named = O.init_grasp_readout(5, bias_std=0.05)
    g = torch.Generator().manual_seed(b * 1000 + n_poses)
    n = n_poses * 42
    n_pad = (n + 127) // 128 * 128
    taps = torch.randn(4, b, n_pad, 128, generator=g) * 0.7
    gy = torch.randn(b, n_poses, generator=g)

    # oracle (fp64) on the valid rows: with every matmul operand rounded to bf16 where the kernels store operand
    # tiles (asserted), and in the reference's plain fp32 arithmetic (printed)
    emb = [taps[s, :, :n].double().reshape(b, n_poses, 42, 128).requires_grad_(True) for s in range(4)]
    p64 = {k: v.double() for k, v in named.items()}
    y_ref, o2 = O.grasp_readout(emb, p64, activation=act, quant='bf16', return_pre_relu=True)
    g_ref = torch.autograd.grad((y_ref * gy.double()).sum(), emb)
    y_ref32 = O.grasp_readout([e.detach().float() for e in emb], named, activation=act)

    pack = ops.grasp_head_pack(_flat(named, O).to(cuda))
    assert pack.numel() == _abi.HEAD_PACK_BYTES
    y, st = ops.grasp_head_fwd(taps.reshape(4, b * n_pad, 128).contiguous().to(cuda), b, n_pad, n_poses, pack,
                               act == 'elu', True)
    e_out, e32 = rel_err(y, y_ref), rel_err(y, y_ref32)
    print(f'[head {act} b={b} np={n_poses}] success vs fp64 oracle {e_out:.2e} (fp32 oracle {e32:.2e})')
    assert e_out < 1e-2
    dtap = ops.grasp_head_bwd(gy.to(cuda), st)
    clear = _unflipped_poses(st, o2.reshape(b * n_poses, 64), f'head {act} b={b} np={n_poses}').reshape(b, n_poses)
    rows_ok = clear[:, :, None].expand(b, n_poses, 42).reshape(b, n)
    for s in range(4):
        got = layout.tiles_to_matrix(dtap[s], 128).view(b, n_pad, 128)
        ref = g_ref[s].reshape(b, n, 128)
        med, p90, mx, l2 = _row_stats(got[:, :n].cpu()[rows_ok], ref[rows_ok])
        print(f'[head {act}] d tap {s}: median {med:.2e} p90 {p90:.2e} max {mx:.2e} L2 {l2:.2e}  '
              f'(all poses: {_row_stats(got[:, :n], ref)})')
        assert p90 < 2e-2 and l2 < 2e-2 and mx < 5e-2, (s, med, p90, mx, l2)
        assert float(got[:, n:].abs().max()) == 0.0 if n_pad > n else True      # padding rows carry no gradient


def test_tile_add_tiles(cuda):
    from clipnerf_b200 import layout, ops
    g = torch.Generator().manual_seed(0)
    a, c = torch.randn(256, 128, generator=g), torch.randn(256, 128, generator=g)
    ta, tc = layout.matrix_to_tiles(a.to(cuda)), layout.matrix_to_tiles(c.to(cuda))
    ops.tile_add_tiles(ta, tc)
    want = (a.bfloat16().float() + c.bfloat16().float()).bfloat16().float()
    assert torch.equal(layout.tiles_to_matrix(ta, 128).cpu(), want)


@pytest.mark.parametrize('rep', ['quaternion', '6d'])
def test_frozen_field_call_uses_the_fused_head(cuda, rep):
    """DeltaNGF._call with the field and the head frozen (DNGFOptimizer's setting, grasp_optimizer.py:158-184):
    embedding + GraspReadout on the CUDA kernels, d(sum success) / d(pose variables) against the oracle and against
    the torch-head path of the same model."""
    from clipnerf_b200 import _abi, params
    from src.lib.delta_ngf.model import DeltaNGF
    from oracle import nerf_oracle as O
    b, nv, n_pts = 2, 3, 48
    sc = O.synthetic_scene(4, b, nv, 8, 48, 64)
    named = O.init_params(9, bias_std=0.02)
    named.update(O.init_grasp_readout(2, bias_std=0.02))
    _, _, img, K, E = sc['inputs']
    t, r = O.synthetic_poses(1, b, n_pts, rep)
    to, ro = t.clone().requires_grad_(True), r.clone().requires_grad_(True)
    ref, o2 = O.query_call(img, K, E, O.compute_matrices(to, ro, rep), sc['features'], named, quant='bf16',
                           head_quant='bf16', return_pre_relu=True)
    ref.sum().backward()

    model = DeltaNGF(n_views=nv, batch_size=b, n_points_train=n_pts, rotation_representation=rep, device=cuda)
    with torch.no_grad():
        model.fine_embedding.flat.copy_(params.from_named(named, 'fine')[:model.fine_embedding.flat.numel()].to(cuda))
    model.grasp_readout.load_keras_variables(named)
    model.set_pose(t, r)
    inputs = (None, None, None, None, img, K, E)
    feats = sc['features'].to(cuda)

    def run():
        y = model._call(inputs, model.compute_matrices(), n_pts, feats)
        return (y,) + torch.autograd.grad(y.sum(), model.pose_variables)
    y_t, gt_t, gr_t = run()                                  # head trainable: torch head
    for p in list(model.grasp_readout.parameters()) + [model.fine_embedding.flat]:
        p.requires_grad_(False)
    names = []
    real = _abi.call

    def spy(name, *a, **k):
        names.append(name)
        return real(name, *a, **k)
    import clipnerf_b200.ops as ops_mod
    states = []
    real_fwd = ops_mod.grasp_head_fwd

    def fwd_spy(*a, **k):
        out = real_fwd(*a, **k)
        states.append(out[1])
        return out
    ops_mod.call, ops_mod.grasp_head_fwd = spy, fwd_spy
    try:
        y_f, gt_f, gr_f = run()
    finally:
        ops_mod.call, ops_mod.grasp_head_fwd = real, real_fwd
    clear = _unflipped_poses(states[0], o2.reshape(b * n_pts, 64), f'fused {rep}').reshape(b, n_pts)
    assert 'cn_grasp_head_fwd' in names and 'cn_grasp_head_bwd' in names, names
    print(f'[fused {rep}] success vs oracle {rel_err(y_f, ref):.2e}, vs torch head {rel_err(y_f, y_t):.2e}')
    assert rel_err(y_f, ref) < 1e-2 and rel_err(y_f, y_t) < 1e-2
    for name, got, want in (('translations', gt_f, to.grad), ('rotations', gr_f, ro.grad)):
        med, p90, mx, l2 = _row_stats(got.cpu()[clear], want[clear])
        print(f'[fused {rep}] grad {name}: median {med:.2e} p90 {p90:.2e} max {mx:.2e} L2 {l2:.2e}; '
              f'vs torch head L2 {_row_stats(got, gt_t if name == "translations" else gr_t)[3]:.2e}')
        assert p90 < 2e-2 and l2 < 5e-2, (name, med, p90, mx, l2)


def test_pose_points_kernels_match_the_broadcast_form(cuda):
    """cn_pose_points_fwd / _bwd against the torch broadcast algebra (the CPU branch of pose_query_points):
    forward, backward and the backward's own backward (the double backward of DeltaNGF.train_step)."""
    from clipnerf_b200.query import pose_query_points
    from src.lib.delta_ngf.model import build_transforms_to_check
    g = torch.Generator().manual_seed(3)
    ttc = build_transforms_to_check(7)
    M = torch.randn(2, 37, 4, 4, generator=g)
    M[..., 3, :] = torch.tensor([0.0, 0.0, 0.0, 1.0])
    wp, wd = torch.randn(2, 37 * 42, 3, generator=g), torch.randn(2, 37 * 42, 3, generator=g)
    v = torch.randn(2, 37, 4, 4, generator=g)

    def run(dev):
        m = M.to(dev).requires_grad_(True)
        a, c = wp.to(dev).requires_grad_(True), wd.to(dev).requires_grad_(True)
        pts, dirs = pose_query_points(m, ttc.to(dev))
        gm, = torch.autograd.grad((pts * a).sum() + (dirs * c).sum(), m, create_graph=True)
        ga, gc = torch.autograd.grad((gm * v.to(dev)).sum(), [a, c])          # d/d(weights) of <dM, v>
        return [t.detach().cpu() for t in (pts, dirs, gm, ga, gc)]
    ref, got = run(torch.device('cpu')), run(cuda)
    assert rel_err(got[0], ref[0]) < 1e-6 and rel_err(got[1], ref[1]) < 1e-6      # (the kernel contracts to FMAs)
    for name, a, c in zip(('dM', 'second-order d_pts', 'second-order d_dirs'), got[2:], ref[2:]):
        if name == 'dM':
            a, c = a[..., :3, :], c[..., :3, :]
        assert rel_err(a, c) < 1e-5, (name, rel_err(a, c))


def test_head_is_deterministic_and_reads_no_uninitialised_memory(cuda, monkeypatch):
    """Five runs of the head forward / backward are bit-identical (no races: every element has one writer), and the
    results do not change when every caller-allocated buffer starts out as NaN patterns instead of whatever the
    allocator hands back -- the padding poses of the last pose tile and the padding rows of a scene are computed from
    unwritten slabs BY DESIGN, and must stay confined to their own (dropped) rows."""
    import clipnerf_b200.ops as ops_mod
    from clipnerf_b200 import _abi
    from oracle import nerf_oracle as O
    named = O.init_grasp_readout(5, bias_std=0.05)
    b, n_poses = 2, 77                                     # 154 poses: 102 padding poses in the second pose tile
    n = n_poses * 42
    n_pad = (n + 127) // 128 * 128
    g = torch.Generator().manual_seed(4)
    taps = (torch.randn(4, b * n_pad, 128, generator=g) * 0.7).to(cuda)
    gy = torch.randn(b, n_poses, generator=g).to(cuda)
    pack = ops_mod.grasp_head_pack(_flat(named, O).to(cuda))
    tiles = torch.stack([ops_mod.tile_add_rows(None, taps[s]) for s in range(4)])

    def run():
        y, st = ops_mod.grasp_head_fwd(tiles, b, n_pad, n_poses, pack, True, True)
        return y.clone(), ops_mod.grasp_head_bwd(gy, st).clone()
    y0, d0 = run()
    for _ in range(4):
        y1, d1 = run()
        assert torch.equal(y0, y1) and torch.equal(d0, d1)
    real_bytes, real_empty = ops_mod._bytes, torch.empty

    def nan_bytes(nbytes, device):
        return torch.full((nbytes,), 0xFF, dtype=torch.uint8, device=device)          # bf16 / fp32 NaN patterns

    def nan_empty(*a, **k):
        t = real_empty(*a, **k)
        if t.is_cuda and t.dtype == torch.uint8:
            t.fill_(0xFF)
        elif t.is_cuda and t.is_floating_point():
            t.fill_(float('nan'))
        return t
    monkeypatch.setattr(ops_mod, '_bytes', nan_bytes)
    monkeypatch.setattr(torch, 'empty', nan_empty)
    y2, d2 = run()
    monkeypatch.setattr(torch, 'empty', real_empty)
    monkeypatch.setattr(ops_mod, '_bytes', real_bytes)
    assert torch.isfinite(y2).all() and torch.equal(y0, y2)
    assert torch.equal(d0, d2)
    assert _abi.LAUNCHES > 0
