"""N2: the encoder's final ``UpSampling2D(2, 'bilinear')`` (layers.py:589,659; legacy_layers.py:190;
model_without.py:76) folded into the gather.  ``CnScene.feat_up == 2`` takes the [H/2, W/2, 256] map the upsample
would have consumed; every test compares with the reference's two-stage form -- ``interpolate_bilinear`` on
``upsample2x_bilinear(map)`` -- evaluated by the oracle (tests/test_oracle.py proves the folded stencil equal to it
on the CPU)."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu

NEAR, FAR = 0.3, 1.3


def _scene(seed, b, nv, nr, H, W, feat_seed=0):
    """synthetic scene whose feature map is [b, nv, H/2, W/2, 256]; 'features_up' is the reference's upsampled map"""
    from oracle import nerf_oracle as O
    sc = O.synthetic_scene(seed, b, nv, nr, H, W)
    g = torch.Generator().manual_seed(1000 + feat_seed)
    sc['features_lr'] = torch.randn(b, nv, H // 2, W // 2, 256, generator=g) * 0.5
    sc['features_up'] = O.upsample2x_bilinear(sc['features_lr'])
    return sc


def _cuda_scenes(sc, cuda):
    from clipnerf_b200 import ops
    ro, rd, img, K, E = [t.to(cuda) for t in sc['inputs']]
    lr = ops.Scene(img, sc['features_lr'].to(cuda), K, E)
    up = ops.Scene(img, sc['features_up'].to(cuda), K, E)
    assert lr.feat_up == 2 and up.feat_up == 1
    return lr, up, ro, rd


def _x_reference(sc, feats, z):
    """fp64 MLP-input rows of the oracle for a feature map `feats` ([b,nv,H,W,256], requires_grad or not)"""
    from oracle import nerf_oracle as O
    i_ro, i_rd, img, K, E = [t.double() for t in sc['inputs']]
    pts = i_ro[:, :, None, :] + z[..., None] * i_rd[:, :, None, :]
    pix_r, cam_r = O.compute_pixel_in_image_mv(pts, K, E)
    g_r = O.get_projection_features_mv(img * 2 - 1.0, feats, pix_r)
    dirs = O.world_to_camera_direction_vector_mv(i_rd, E)[:, :, :, None, :].expand(-1, -1, -1, z.shape[-1], -1)
    return torch.cat([O.position_encoding(cam_r[..., :3]), O.position_encoding(dirs), g_r], -1).reshape(-1, 379)


def _dx_tiles(gx, cuda):
    from clipnerf_b200 import layout
    cols = torch.tensor([layout.x_col_to_ref(j) for j in range(379)])
    gint = torch.zeros(gx.shape[0], 384)
    gint[:, :379] = gx[:, cols]
    return layout.matrix_to_tiles(gint).to(cuda)


@pytest.mark.parametrize('shape', [(2, 3, 16, 24, 32), (1, 2, 8, 4, 6), (1, 1, 8, 2, 2)])
def test_debug_gather_is_the_two_stage_form_bit_exact(cuda, shape):
    """cn_project_gather_debug with feat_up == 2 upsamples texel by texel in the reference's own operation order:
    bit-exact against interpolate_bilinear(upsample2x_bilinear(x)) of the oracle, down to 1x1 low-resolution maps"""
    from clipnerf_b200 import ops
    from oracle import nerf_oracle as O
    b, nv, nr, H, W = shape
    sc = _scene(1, b, nv, nr, H, W)
    lr, up, ro, rd = _cuda_scenes(sc, cuda)
    i_ro, i_rd, img, K, E = sc['inputs']
    pts, z = O.sample_along_ray(i_ro, i_rd, NEAR, FAR, 64, sc['u_strat'])
    pix_r, _ = O.compute_pixel_in_image_mv(pts, K, E)
    g_r = O.get_projection_features_mv(img * 2 - 1.0, sc['features_up'], pix_r)
    pix, corner, gth = ops.project_gather_debug(lr, ro, rd, z.to(cuda))
    assert torch.equal(pix.cpu(), pix_r)
    assert torch.equal(gth.cpu(), g_r)
    _, corner_up, gth_up = ops.project_gather_debug(up, ro, rd, z.to(cuda))
    assert torch.equal(corner, corner_up) and torch.equal(gth, gth_up)


@pytest.mark.parametrize('shape', [(1, 3, 64, 24, 32), (2, 2, 32, 42, 76), (1, 2, 128, 4, 6), (1, 1, 64, 2, 2)])
def test_folded_gather_rows_against_the_two_stage_gather(cuda, shape):
    """MLP-input rows of the binned low-resolution gather: the feature columns are the bf16 rounding of a value within
    fp32 rounding (1e-6 of the map's range) of the two-stage fp32 gather, every other column is bit-identical.  Ragged
    tiles (21 x 38 low-resolution cells), work items chunked at 256 rows, clamped out-of-frustum samples."""
    from clipnerf_b200 import layout, ops
    b, nv, nr, H, W = shape
    sc = _scene(5, b, nv, nr, H, W)
    lr, up, ro, rd = _cuda_scenes(sc, cuda)
    z = ops.sample_coarse(sc['u_strat'].to(cuda), NEAR, FAR)
    _, _, gth = ops.project_gather_debug(lr, ro, rd, z)
    want = gth.reshape(-1, 259)
    got = layout.x_matrix_to_ref(layout.tiles_to_matrix(ops.project_gather_fwd(lr, ro, rd, z), 384))
    ref = layout.x_matrix_to_ref(layout.tiles_to_matrix(ops.project_gather_fwd(up, ro, rd, z), 384))
    assert torch.equal(got[:, :123], ref[:, :123])               # posenc(p), posenc(d), image columns
    scale = float(sc['features_lr'].abs().max())
    # a bf16 value is within half an ulp (2^-9 relative) of the fp32 value it rounds; the folded stencil and the
    # two-stage form differ by a few fp32 ulps before that rounding
    err = (got[:, 123:] - want[:, 3:]).abs()
    bound = want[:, 3:].abs() * 2.0 ** -8 + 4e-6 * scale
    assert bool((err <= bound).all()), float((err - bound).max())
    same = float((got[:, 123:] == want[:, 3:].to(torch.bfloat16).float()).float().mean())
    assert same > 0.995, same                                    # rounding flips only


@pytest.mark.parametrize('shape,corner', [((1, 2, 4, 24, 32), False), ((1, 2, 64, 24, 32), False),
                                          ((2, 2, 32, 42, 76), False), ((1, 2, 192, 24, 32), True)])
def test_folded_gather_backward_against_fp64_autograd(cuda, shape, corner):
    """d_features (gradient w.r.t. the LOW-resolution map) and dz of cn_project_gather_bwd against fp64 autograd through
    upsample2x_bilinear -> interpolate_bilinear; corner=True clamps every sample onto one frame corner (24 576 rows
    on one cell: the split path of the sweep); accumulate / write-once / workspace re-use as for the full-size map"""
    from clipnerf_b200 import ops
    from oracle import nerf_oracle as O
    b, nv, nr, H, W = shape
    sc = _scene(6, b, nv, nr, H, W)
    if corner:
        K = sc['inputs'][3].clone()
        K[..., 0, 2] = -4000.0
        K[..., 1, 2] = 5000.0
        sc['inputs'] = sc['inputs'][:3] + (K,) + sc['inputs'][4:]
    lr, _, ro, rd = _cuda_scenes(sc, cuda)
    feats = sc['features_lr'].double().requires_grad_(True)
    _, z32 = O.sample_along_ray(sc['inputs'][0], sc['inputs'][1], NEAR, FAR, 64, sc['u_strat'])
    z = z32.double().requires_grad_(True)
    x_ref = _x_reference(sc, O.upsample2x_bilinear(feats), z)
    gx = torch.randn(x_ref.shape[0], 379, generator=torch.Generator().manual_seed(10))
    gx[:, :60] *= 1e-3
    gx = gx.to(torch.bfloat16).to(torch.float32)
    (x_ref * gx.double()).sum().backward()
    dx = _dx_tiles(gx, cuda)
    x, ws = ops.project_gather_fwd(lr, ro, rd, z32.to(cuda), return_ws=True)
    d_feat = torch.zeros_like(lr.features)
    assert tuple(d_feat.shape) == (b, nv, H // 2, W // 2, 256)
    dz = torch.zeros(b, nr, 64, device=cuda)
    ops.project_gather_bwd(lr, ro, rd, z32.to(cuda), dx, d_feat, dz, ws)
    assert rel_err(d_feat, feats.grad) < 1e-4, rel_err(d_feat, feats.grad)
    assert rel_err(dz, z.grad) < 2e-3, rel_err(dz, z.grad)
    d_once = torch.full_like(d_feat, float('nan'))               # accumulate = 0 starts from zero itself
    ops.feature_grad(lr, [(ws, dx, nr, 64)], d_once, accumulate=False)
    assert rel_err(d_once, feats.grad) < 1e-4
    ops.project_gather_bwd(lr, ro, rd, z32.to(cuda), dx, d_feat, dz, ws)
    assert rel_err(d_feat, 2 * feats.grad) < 1e-4


def test_folded_feature_grad_of_two_row_sets_in_one_sweep(cuda):
    """coarse (S = 64) + fine-sized (S = 128) rows summed by one sweep == the sum of two single-set sweeps; the
    per-channel checksum: every stencil's weights sum to one, so sum over texels == sum over rows of dX"""
    from clipnerf_b200 import layout, ops
    sc = _scene(7, 1, 3, 96, 40, 56)
    lr, _, ro, rd = _cuda_scenes(sc, cuda)
    g = torch.Generator(device=cuda).manual_seed(4)
    sets, want = [], torch.zeros(3, 256, dtype=torch.float64, device=cuda)
    for S in (64, 128):
        u = torch.rand(1, 96, S, generator=g, device=cuda)
        z = (NEAR + (FAR - NEAR) * (torch.arange(S, device=cuda) + u) / S).contiguous()
        _x, ws = ops.project_gather_fwd(lr, ro, rd, z, return_ws=True)
        rows = 3 * 96 * S
        gx = torch.zeros(rows, 384, device=cuda)
        gx[:, :256] = torch.randn(rows, 256, generator=g, device=cuda)
        gx = gx.to(torch.bfloat16).to(torch.float32)
        want += gx[:, :256].view(3, 96 * S, 256).double().sum(1)
        sets.append((ws, layout.matrix_to_tiles(gx), 96, S))
    both = torch.full_like(lr.features, float('nan'))
    ops.feature_grad(lr, sets, both, accumulate=False)
    one = torch.zeros_like(both)
    ops.feature_grad(lr, sets[:1], one, accumulate=True)
    ops.feature_grad(lr, sets[1:], one, accumulate=True)
    assert rel_err(both, one) < 1e-5, rel_err(both, one)
    assert rel_err(both[0].double().sum(dim=(1, 2)), want) < 1e-5


@pytest.mark.parametrize('cap', [0, 5])
def test_folded_feature_grad_with_a_full_split_list(cuda, cap, monkeypatch):
    """CN_DEBUG_HEAVY_CAP forces the list-full path of the low-resolution sweep (cell rows that do not fit are summed
    in place, their reserved slots become skip entries)"""
    from clipnerf_b200 import ops
    sc = _scene(8, 1, 2, 512, 12, 16)
    lr, _, ro, rd = _cuda_scenes(sc, cuda)
    z = ops.sample_coarse(sc['u_strat'].to(cuda), NEAR, FAR)
    x, ws = ops.project_gather_fwd(lr, ro, rd, z, return_ws=True)
    dx = (torch.randn(x.numel() // 2, generator=torch.Generator().manual_seed(1)) * 0.5).to(torch.bfloat16).view(
        torch.uint8).to(cuda)
    ref = torch.full_like(lr.features, float('nan'))
    ops.feature_grad(lr, [(ws, dx, 512, 64)], ref, accumulate=False)
    assert torch.isfinite(ref).all()
    monkeypatch.setenv('CN_DEBUG_HEAVY_CAP', str(cap))
    got = torch.full_like(ref, float('nan'))
    ops.feature_grad(lr, [(ws, dx, 512, 64)], got, accumulate=False)
    assert rel_err(got, ref) < 1e-5, rel_err(got, ref)
    ops.feature_grad(lr, [(ws, dx, 512, 64)], got, accumulate=True)
    assert rel_err(got, 2 * ref) < 1e-5


def test_render_with_lowres_features_against_the_oracle(cuda):
    """the whole render call on the pre-upsample map: outputs and the coarse-pass gradients (z is identical on both
    sides) against the oracle's upsample -> _call, bf16-operand form; the fine outputs with the end-to-end bars"""
    from clipnerf_b200 import params, render
    from oracle import nerf_oracle as O
    sc = _scene(11, 1, 3, 32, 48, 64)
    named = O.init_params(5, bias_std=0.02)
    p = {k: v.clone().requires_grad_(True) for k, v in named.items()}
    feats = sc['features_lr'].clone().requires_grad_(True)
    ref = O.mvnerf_call(sc['inputs'], feats, p, near=NEAR, far=FAR, n_samples=64, u_strat=sc['u_strat'],
                        u_pdf=sc['u_pdf'], quant='bf16', feat_up=2)
    (O.mse(sc['target'], ref[0]) + 0.1 * ref[1].mean()).backward()
    pc = params.from_named(named, 'coarse').to(cuda).requires_grad_(True)
    pf = params.from_named(named, 'fine').to(cuda).requires_grad_(True)
    fd = sc['features_lr'].to(cuda).requires_grad_(True)
    out = render.render_rays(pc, pf, [t.to(cuda) for t in sc['inputs']], fd, sc['u_strat'].to(cuda),
                             sc['u_pdf'].to(cuda), NEAR, FAR)
    tgt = sc['target'].to(cuda)
    (torch.mean((out[0] - tgt) ** 2) + 0.1 * out[1].mean()).backward()
    errs = {n: rel_err(a, r) for n, a, r in zip(('rgb', 'depth', 'fine_rgb', 'fine_depth'), out, ref)}
    for name, gv in params.to_named(pc.grad.cpu(), 'coarse').items():
        errs['grad ' + name] = rel_err(gv, p[name].grad)
    assert tuple(fd.grad.shape) == tuple(sc['features_lr'].shape)
    errs['grad features_lr'] = rel_err(fd.grad, feats.grad)
    for k, v in errs.items():
        print(f'[lowres render] {k:62s} {v:.3e}')
    bad = {k: v for k, v in errs.items() if not v < (2e-2 if k.startswith('grad') else 1.5e-2 if k.startswith('fine') else 1e-2)}
    assert not bad, bad


def test_lowres_scene_rejects_odd_image_sizes(cuda):
    from clipnerf_b200 import ops
    from clipnerf_b200._abi import CnScene
    import ctypes
    sc = _scene(1, 1, 1, 2, 24, 32)
    lr, _, ro, rd = _cuda_scenes(sc, cuda)
    bad = CnScene(lr.images.data_ptr(), lr.features.data_ptr(), lr.intrinsics.data_ptr(), lr.extr_inv.data_ptr(),
                  1, 1, 23, 32, 2)
    z = ops.sample_coarse(sc['u_strat'].to(cuda), NEAR, FAR)
    ws = ops.gather_workspace(lr, 2, 64)
    with pytest.raises(RuntimeError, match='feat_up'):
        ops.call('cn_project_bin', ctypes.byref(bad), ro.data_ptr(), rd.data_ptr(), z.data_ptr(), 2, 64, ws.data_ptr(),
                 ws.numel(), None)
