"""Per-kernel parity of the SIMT kernels (sampling, projection + gather, compositing) against the
CPU oracle, through the C ABI.  Bit-exact where the path is index / bookkeeping work, fp tolerance
(stated per test) elsewhere."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu

NEAR, FAR = 0.3, 1.3


def _scene(seed=0, b=1, nv=3, nr=8, H=24, W=32):
    from oracle import nerf_oracle as O
    return O.synthetic_scene(seed, b, nv, nr, H, W)


def _cuda_scene(sc, cuda):
    from clipnerf_b200 import ops
    ro, rd, img, K, E = [t.to(cuda) for t in sc['inputs']]
    return ops.Scene(img, sc['features'].to(cuda), K, E), ro, rd


def test_sample_coarse_bit_exact(cuda):
    from clipnerf_b200 import ops
    from oracle import nerf_oracle as O
    sc = _scene(nr=64)
    ro, rd = sc['inputs'][0], sc['inputs'][1]
    _, z_ref = O.sample_along_ray(ro, rd, NEAR, FAR, 64, sc['u_strat'])
    z = ops.sample_coarse(sc['u_strat'].to(cuda), NEAR, FAR)
    assert torch.equal(z.cpu(), z_ref)


@pytest.mark.parametrize('peaky', [False, True])
def test_sample_pdf_fwd_bit_exact(cuda, peaky):
    """bins (above), merged z values and the stable sort permutation are bit-exact given (z, w, u)."""
    from clipnerf_b200 import ops
    from oracle import nerf_oracle as O
    g = torch.Generator().manual_seed(3)
    n = 512
    u_s = torch.rand(1, n, 64, generator=g)
    z = torch.tensor([NEAR + i * (FAR - NEAR) / 64 for i in range(64)], dtype=torch.float32) + u_s * ((FAR - NEAR) / 64)
    w = torch.rand(1, n, 64, generator=g)
    if peaky:
        w = w * 1e-6
        w[..., 20] = 0.9                       # single spike: every draw lands next to bin 19/20
        w[0, :8] = 0.0                         # all-zero weights: uniform pdf from the 1e-5 floor
    u = torch.rand(1, n, 64, generator=g)
    u[0, 0, :4] = torch.tensor([0.0, 0.99999994, 0.5, 1e-8])   # edges of the cdf (quirk Q5)
    z_mid = 0.5 * (z[..., 1:] + z[..., :-1])
    zs_ref, above_ref, _ = O.sample_pdf(z_mid, w[..., 1:-1], 64, u, return_indices=True)
    all_ref, perm_ref = torch.sort(torch.cat([z, zs_ref], -1), dim=-1, stable=True)
    z_fine, perm, above = ops.sample_pdf_fwd(z.to(cuda), w.to(cuda), u.to(cuda))
    assert torch.equal(above.cpu().to(torch.int32), above_ref)
    assert torch.equal(z_fine.cpu(), all_ref)
    assert torch.equal(perm.cpu().to(torch.int64), perm_ref)


def test_sample_pdf_bwd(cuda):
    from clipnerf_b200 import ops
    from oracle import nerf_oracle as O
    g = torch.Generator().manual_seed(4)
    n = 256
    z = torch.tensor([NEAR + i * (FAR - NEAR) / 64 for i in range(64)], dtype=torch.float32) + \
        torch.rand(1, n, 64, generator=g) * ((FAR - NEAR) / 64)
    w = torch.rand(1, n, 64, generator=g) * 0.1
    u = torch.rand(1, n, 64, generator=g)
    gz = torch.randn(1, n, 128, generator=g)
    w64 = w.double().requires_grad_(True)
    z64 = z.double()
    zs = O.sample_pdf(0.5 * (z64[..., 1:] + z64[..., :-1]), w64[..., 1:-1], 64, u.double())
    allz, _ = torch.sort(torch.cat([z64, zs], -1), dim=-1, stable=True)
    (allz * gz.double()).sum().backward()
    z_fine, perm, _ = ops.sample_pdf_fwd(z.to(cuda), w.to(cuda), u.to(cuda))
    dw = ops.sample_pdf_bwd(z.to(cuda), w.to(cuda), u.to(cuda), perm, gz.to(cuda))
    e = rel_err(dw, w64.grad)
    assert e < 1e-2, e      # fp32 kernel vs fp64 autograd; 1/denom^2 terms cancel heavily


@pytest.mark.parametrize('S', [64, 128])
def test_composite_fwd_bwd(cuda, S):
    """rel-err <= 1e-4 (fp32 scan order differs from the sequential cumprod)."""
    from clipnerf_b200 import ops
    from oracle import nerf_oracle as O
    g = torch.Generator().manual_seed(5)
    n = 300
    z = torch.sort(torch.rand(1, n, S, generator=g) * (FAR - NEAR) + NEAR, dim=-1).values
    rgbs = torch.rand(1, n, S, 4, generator=g)
    rgbs[..., 3] = rgbs[..., 3] * 30.0            # densities up to 30: transmittance really decays
    rgbs[0, 0, :, 3] = 5.0                         # uniform sigma: closed form below
    d_rgb = torch.randn(1, n, 3, generator=g)
    d_depth = torch.randn(1, n, generator=g)
    d_w = torch.randn(1, n, S, generator=g)

    z64 = z.double().requires_grad_(True)
    c64 = rgbs[..., :3].double().requires_grad_(True)
    s64 = rgbs[..., 3].double().requires_grad_(True)
    rgb_r, depth_r, w_r = O.volumetric_render(z64, s64, c64)
    ((rgb_r * d_rgb.double()).sum() + (depth_r * d_depth.double()).sum() + (w_r * d_w.double()).sum()).backward()

    rgb, depth, w = ops.composite_fwd(z.to(cuda), rgbs.to(cuda))
    assert rel_err(rgb, rgb_r.detach()) < 1e-4
    assert rel_err(depth, depth_r.detach()) < 1e-4
    assert rel_err(w, w_r.detach()) < 1e-4
    # closed form for uniform sigma: T_i = exp(-sigma (z_i - z_0))
    T = torch.exp(-5.0 * (z[0, 0] - z[0, 0, 0])).double()
    a = 1 - torch.exp(-5.0 * torch.cat([z[0, 0, 1:] - z[0, 0, :-1], z[0, 0, -1:] - z[0, 0, -2:-1]])).double()
    assert rel_err(w[0, 0], a * T) < 1e-4

    d_y, dz = ops.composite_bwd(z.to(cuda), rgbs.to(cuda), d_rgb.to(cuda), d_depth.to(cuda), d_w.to(cuda), True)
    c = rgbs[..., :3].double()
    dy_ref = torch.cat([c64.grad * c * (1 - c), (s64.grad * (1 - torch.exp(-rgbs[..., 3].double())))[..., None]], -1)
    assert rel_err(d_y, dy_ref) < 1e-3, rel_err(d_y, dy_ref)
    assert rel_err(dz, z64.grad) < 1e-3, rel_err(dz, z64.grad)


def test_project_gather_debug_bit_exact(cuda):
    """pixel coordinates, corner indices and the fp32 bilinear gather are bit-exact given the same z."""
    from clipnerf_b200 import ops
    from oracle import nerf_oracle as O
    sc = _scene(seed=1, b=2, nv=3, nr=16)
    scene, ro, rd = _cuda_scene(sc, cuda)
    i_ro, i_rd, img, K, E = sc['inputs']
    pts, z = O.sample_along_ray(i_ro, i_rd, NEAR, FAR, 64, sc['u_strat'])
    pix_r, cam_r = O.compute_pixel_in_image_mv(pts, K, E)
    g_r = O.get_projection_features_mv(img * 2 - 1.0, sc['features'], pix_r)
    pix, corner, gth = ops.project_gather_debug(scene, ro, rd, z.to(cuda))
    assert torch.equal(pix.cpu(), pix_r)
    H, W = img.shape[2], img.shape[3]
    fx = torch.clamp(torch.floor(pix_r[..., 0]), 0, W - 2).to(torch.int32)
    fy = torch.clamp(torch.floor(pix_r[..., 1]), 0, H - 2).to(torch.int32)
    assert torch.equal(corner.cpu()[..., 0], fx) and torch.equal(corner.cpu()[..., 1], fy)
    assert torch.equal(gth.cpu(), g_r)


def test_project_gather_fwd_tiles(cuda):
    """bf16 MLP-input rows: rel-err <= 4e-3 (one bf16 rounding) against the oracle's fp32 concat."""
    from clipnerf_b200 import layout, ops
    from oracle import nerf_oracle as O
    sc = _scene(seed=2, b=2, nv=3, nr=16)
    scene, ro, rd = _cuda_scene(sc, cuda)
    i_ro, i_rd, img, K, E = sc['inputs']
    pts, z = O.sample_along_ray(i_ro, i_rd, NEAR, FAR, 64, sc['u_strat'])
    pix_r, cam_r = O.compute_pixel_in_image_mv(pts, K, E)
    g_r = O.get_projection_features_mv(img * 2 - 1.0, sc['features'], pix_r)
    dirs = O.world_to_camera_direction_vector_mv(i_rd, E)[:, :, :, None, :].expand(-1, -1, -1, 64, -1)
    x_ref = torch.cat([O.position_encoding(cam_r[..., :3]), O.position_encoding(dirs), g_r], -1).reshape(-1, 379)
    x = ops.project_gather_fwd(scene, ro, rd, z.to(cuda))
    xm = layout.tiles_to_matrix(x, 384)
    assert float(xm[:, 379:].abs().max()) == 0.0
    got = layout.x_matrix_to_ref(xm).cpu()
    # posenc arguments reach ~2e3 rad at the top frequency: sin/cos differ by the argument's fp32
    # rounding between libm implementations, so compare absolutely on [-1, 1] values
    assert float((got[:, :120] - x_ref[:, :120]).abs().max()) < 1e-2
    assert rel_err(got[:, 120:], x_ref[:, 120:]) < 4e-3


@pytest.mark.parametrize('need_feat', [True, False])
def test_project_gather_bwd(cuda, need_feat):
    from clipnerf_b200 import layout, ops
    from oracle import nerf_oracle as O
    sc = _scene(seed=3, b=1, nv=2, nr=4)
    scene, ro, rd = _cuda_scene(sc, cuda)
    i_ro, i_rd, img, K, E = [t.double() for t in sc['inputs']]
    feats = sc['features'].double().requires_grad_(True)
    _, z32 = O.sample_along_ray(sc['inputs'][0], sc['inputs'][1], NEAR, FAR, 64, sc['u_strat'])
    z = z32.double().requires_grad_(True)
    pts = i_ro[:, :, None, :] + z[..., None] * i_rd[:, :, None, :]
    pix_r, cam_r = O.compute_pixel_in_image_mv(pts, K, E)
    g_r = O.get_projection_features_mv(img * 2 - 1.0, feats, pix_r)
    dirs = O.world_to_camera_direction_vector_mv(i_rd, E)[:, :, :, None, :].expand(-1, -1, -1, 64, -1)
    x_ref = torch.cat([O.position_encoding(cam_r[..., :3]), O.position_encoding(dirs), g_r], -1).reshape(-1, 379)
    gen = torch.Generator().manual_seed(9)
    gx = torch.randn(x_ref.shape[0], 379, generator=gen).to(torch.bfloat16).to(torch.float32)
    gx[:, :60] *= 1e-3        # posenc gradients are amplified by f_k up to ~1.6e3
    gx = gx.to(torch.bfloat16).to(torch.float32)
    (x_ref * gx.double()).sum().backward()
    # dX tile images in the internal column order
    cols = torch.tensor([layout.x_col_to_ref(j) for j in range(379)])
    gint = torch.zeros(x_ref.shape[0], 384)
    gint[:, :379] = gx[:, cols]
    dx = layout.matrix_to_tiles(gint).to(cuda)
    d_feat = torch.zeros_like(scene.features) if need_feat else None
    dz = torch.zeros(1, 4, 64, device=cuda)
    ops.project_gather_bwd(scene, ro, rd, z32.to(cuda), dx, d_feat, dz)
    if need_feat:
        assert rel_err(d_feat, feats.grad) < 1e-4, rel_err(d_feat, feats.grad)
    assert rel_err(dz, z.grad) < 2e-3, rel_err(dz, z.grad)


@pytest.mark.parametrize('shape', [(1, 3, 64, 24, 32), (2, 2, 32, 41, 75)])
def test_binned_gather_matches_direct_gather(cuda, shape):
    """many rows per bin (work items are chunked at 128 rows), ragged border tiles (41x75 image), clamped
    out-of-frustum samples: the feature columns of X equal the bf16 rounding of the bit-exact fp32 gather."""
    from clipnerf_b200 import layout, ops
    b, nv, nr, H, W = shape
    sc = _scene(seed=5, b=b, nv=nv, nr=nr, H=H, W=W)
    scene, ro, rd = _cuda_scene(sc, cuda)
    z = ops.sample_coarse(sc['u_strat'].to(cuda), NEAR, FAR)
    _, _, gth = ops.project_gather_debug(scene, ro, rd, z)
    x = ops.project_gather_fwd(scene, ro, rd, z)
    got = layout.x_matrix_to_ref(layout.tiles_to_matrix(x, 384))
    want = gth.reshape(-1, 259)
    assert torch.equal(got[:, 123:], want[:, 3:].to(torch.bfloat16).to(torch.float32))
    assert torch.equal(got[:, 120:123], want[:, :3].to(torch.bfloat16).to(torch.float32))


@pytest.mark.parametrize('corner', [False, True])
def test_binned_scatter_many_rows(cuda, corner):
    """feature-map gradient and dz with hundreds of rows per bin against fp64 autograd of the oracle.
    corner=True moves the principal point far outside the frame, so every sample is clamped to the same
    frame corner: 12288 rows on one cell (the 'huge texel' path that spreads a texel over all CTAs)."""
    from clipnerf_b200 import layout, ops
    from oracle import nerf_oracle as O
    sc = _scene(seed=6, b=1, nv=2, nr=192 if corner else 64, H=24, W=32)
    if corner:
        K = sc['inputs'][3].clone()
        K[..., 0, 2] = -4000.0
        K[..., 1, 2] = 5000.0
        sc['inputs'] = sc['inputs'][:3] + (K,) + sc['inputs'][4:]
    scene, ro, rd = _cuda_scene(sc, cuda)
    i_ro, i_rd, img, K, E = [t.double() for t in sc['inputs']]
    feats = sc['features'].double().requires_grad_(True)
    _, z32 = O.sample_along_ray(sc['inputs'][0], sc['inputs'][1], NEAR, FAR, 64, sc['u_strat'])
    z = z32.double().requires_grad_(True)
    pts = i_ro[:, :, None, :] + z[..., None] * i_rd[:, :, None, :]
    pix_r, cam_r = O.compute_pixel_in_image_mv(pts, K, E)
    g_r = O.get_projection_features_mv(img * 2 - 1.0, feats, pix_r)
    dirs = O.world_to_camera_direction_vector_mv(i_rd, E)[:, :, :, None, :].expand(-1, -1, -1, 64, -1)
    x_ref = torch.cat([O.position_encoding(cam_r[..., :3]), O.position_encoding(dirs), g_r], -1).reshape(-1, 379)
    gen = torch.Generator().manual_seed(10)
    gx = torch.randn(x_ref.shape[0], 379, generator=gen)
    gx[:, :60] *= 1e-3
    gx = gx.to(torch.bfloat16).to(torch.float32)
    (x_ref * gx.double()).sum().backward()
    cols = torch.tensor([layout.x_col_to_ref(j) for j in range(379)])
    gint = torch.zeros(x_ref.shape[0], 384)
    gint[:, :379] = gx[:, cols]
    dx = layout.matrix_to_tiles(gint).to(cuda)
    x, ws = ops.project_gather_fwd(scene, ro, rd, z32.to(cuda), return_ws=True)
    d_feat = torch.zeros_like(scene.features)
    dz = torch.zeros(1, z32.shape[1], 64, device=cuda)
    ops.project_gather_bwd(scene, ro, rd, z32.to(cuda), dx, d_feat, dz, ws)
    assert rel_err(d_feat, feats.grad) < 1e-4, rel_err(d_feat, feats.grad)
    # one sweep that writes every texel exactly once (no zero fill) gives the same map
    d_once = torch.full_like(d_feat, float('nan'))
    ops.feature_grad(scene, [(ws, dx, z32.shape[1], 64)], d_once, accumulate=False)
    assert rel_err(d_once, feats.grad) < 1e-4
    assert rel_err(dz, z.grad) < 2e-3, rel_err(dz, z.grad)
    # the workspace can be re-used: a second backward accumulates the same amounts again
    ops.project_gather_bwd(scene, ro, rd, z32.to(cuda), dx, d_feat, dz, ws)
    assert rel_err(d_feat, 2 * feats.grad) < 1e-4


@pytest.mark.parametrize('cap', [0, 3, 17])
def test_feature_grad_with_a_full_deferred_list(cuda, cap, monkeypatch):
    """many texels with hundreds of rows (small map, many rays): the deferred-chunk list of the feature-gradient sweep
    is sized from the row count and cannot fill up in production; CN_DEBUG_HEAVY_CAP forces the list-full path
    (texels that do not fit are summed in place, the slots they had reserved are skip entries) -- the map must equal
    the one of the default run, with and without accumulation, on a workspace that holds stale entries"""
    from clipnerf_b200 import ops
    sc = _scene(seed=8, b=1, nv=2, nr=512, H=6, W=8)
    scene, ro, rd = _cuda_scene(sc, cuda)
    z = ops.sample_coarse(sc['u_strat'].to(cuda), NEAR, FAR)
    x, ws = ops.project_gather_fwd(scene, ro, rd, z, return_ws=True)
    dx = (torch.randn(x.numel() // 2, generator=torch.Generator().manual_seed(1)) * 0.5).to(torch.bfloat16).view(
        torch.uint8).to(cuda)
    ref = torch.full_like(scene.features, float('nan'))
    ops.feature_grad(scene, [(ws, dx, 512, 64)], ref, accumulate=False)          # also leaves entries in the list
    assert torch.isfinite(ref).all()
    monkeypatch.setenv('CN_DEBUG_HEAVY_CAP', str(cap))
    got = torch.full_like(ref, float('nan'))
    ops.feature_grad(scene, [(ws, dx, 512, 64)], got, accumulate=False)
    assert rel_err(got, ref) < 1e-5, rel_err(got, ref)
    ops.feature_grad(scene, [(ws, dx, 512, 64)], got, accumulate=True)
    assert rel_err(got, 2 * ref) < 1e-5


def test_project_gather_bwd_rays(cuda):
    """cn_project_gather_bwd_rays with S > 1 and z != 0: d/d(ray origin), d/d(ray direction) (through x = o + z d AND
    through posenc(E^-1 [d; 1])) and dz against the fp64 oracle for a given dX."""
    from clipnerf_b200 import layout, ops
    from oracle import nerf_oracle as O
    sc = _scene(seed=5, b=2, nv=2, nr=4)
    scene, ro, rd = _cuda_scene(sc, cuda)
    i_ro, i_rd, img, K, E = [t.double() for t in sc['inputs']]
    i_ro, i_rd = i_ro.requires_grad_(True), i_rd.requires_grad_(True)
    _, z32 = O.sample_along_ray(sc['inputs'][0], sc['inputs'][1], NEAR, FAR, 64, sc['u_strat'])
    z = z32.double().requires_grad_(True)
    pts = i_ro[:, :, None, :] + z[..., None] * i_rd[:, :, None, :]
    pix_r, cam_r = O.compute_pixel_in_image_mv(pts, K, E)
    g_r = O.get_projection_features_mv(img * 2 - 1.0, sc['features'].double(), pix_r)
    dirs = O.world_to_camera_direction_vector_mv(i_rd, E)[:, :, :, None, :].expand(-1, -1, -1, 64, -1)
    x_ref = torch.cat([O.position_encoding(cam_r[..., :3]), O.position_encoding(dirs), g_r], -1).reshape(-1, 379)
    gen = torch.Generator().manual_seed(11)
    gx = torch.randn(x_ref.shape[0], 379, generator=gen)
    gx[:, :120] *= 1e-3       # posenc gradients are amplified by f_k up to ~1.6e3
    gx = gx.to(torch.bfloat16).to(torch.float32)
    (x_ref * gx.double()).sum().backward()
    cols = torch.tensor([layout.x_col_to_ref(j) for j in range(379)])
    gint = torch.zeros(x_ref.shape[0], 384)
    gint[:, :379] = gx[:, cols]
    dx = layout.matrix_to_tiles(gint).to(cuda)
    ws = ops.project_bin(scene, ro, rd, z32.to(cuda))
    d_o, d_d, dz = ops.project_gather_bwd_rays(scene, ro, rd, z32.to(cuda), dx, ws, need_dz=True)
    errs = dict(d_o=rel_err(d_o, i_ro.grad), d_d=rel_err(d_d, i_rd.grad), dz=rel_err(dz, z.grad))
    print('[bwd_rays]', errs)
    assert max(errs.values()) < 2e-3, errs
