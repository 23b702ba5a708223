"""Pins the oracle against the reference's OWN source files, executed on the torch-backed TensorFlow
stand-in (oracle/tf_shim).  Runs only where the reference checkout exists (the build container);
the golden vectors it produces for the other machines are checked in test_golden.py."""
import os

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O
from oracle import tf_shim

pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(tf_shim.REFERENCE_ROOT, 'src')),
                                reason='reference checkout not present on this machine')

NEAR, FAR = 0.3, 1.3


@pytest.fixture(scope='module')
def ref():
    return tf_shim.load_reference()


def test_host_functions_identical(ref):
    nu, _, _ = ref
    pose = O.polar_pose(0.9, 0.6, 0.8)
    K = np.array([[55.4, 0, 32], [0, 55.4, 24], [0, 0, 1]], dtype=np.float32)
    for a, b in zip(nu.get_rays(64, 48, pose, K), O.get_rays(64, 48, pose, K)):
        assert np.array_equal(a, b)
    np.random.seed(3)
    r_ref = nu.bbox_biased_sample(100, np.array([0, 0, 48, 64]), 48, 64)
    r_or = O.bbox_biased_sample(100, np.array([0, 0, 48, 64]), 48, 64, rng=np.random.RandomState(3))
    assert np.array_equal(r_ref, r_or)


@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-11), (torch.float32, 2e-5)])
def test_op_functions_match_reference_source(ref, dtype, tol):
    nu, _, _ = ref
    sc = O.synthetic_scene(31, 2, 3, 6, 20, 28, dtype=dtype)
    ro, rd, img, K, E = sc['inputs']
    b, nr, S = 2, 6, 64
    tf_shim.queue_uniform(sc['u_strat'])
    pts_ref, z_ref = nu.sample_along_ray(ro, rd, NEAR, FAR, b, nr, S)
    pts, z = O.sample_along_ray(ro, rd, NEAR, FAR, S, sc['u_strat'])
    assert torch.allclose(z, z_ref, rtol=0, atol=tol) and torch.allclose(pts, pts_ref, rtol=0, atol=tol)
    pix_ref, cam_ref = nu.compute_pixel_in_image_mv(pts_ref, K, E)
    pix, cam = O.compute_pixel_in_image_mv(pts_ref, K, E)
    assert pix.shape == pix_ref.shape and cam.shape == cam_ref.shape
    assert float((cam - cam_ref).abs().max()) < tol * 10 and float((pix - pix_ref).abs().max() / pix_ref.abs().max()) < tol * 10
    d_ref = nu.world_to_camera_direction_vector_mv(rd, E, 3)
    assert float((O.world_to_camera_direction_vector_mv(rd, E) - d_ref).abs().max()) < tol * 10
    g_ref = nu.get_projection_features_mv(img * 2 - 1.0, sc['features'], pix_ref, nr, S, b)
    assert torch.equal(O.get_projection_features_mv(img * 2 - 1.0, sc['features'], pix_ref), g_ref)
    enc_ref = nu.position_encoding(cam_ref[..., :3].reshape(b * 3, nr, S, 3), 10, np.pi)
    enc = O.position_encoding(cam_ref[..., :3].reshape(b * 3, nr, S, 3), 10, np.pi)
    assert torch.equal(enc, enc_ref)
    sig = torch.rand(b, nr, S, dtype=dtype) * 8
    assert torch.equal(O.sigma_to_alpha(sig, z_ref), nu.sigma_to_alpha(sig, z_ref))
    w = torch.rand(b, nr, S - 2, dtype=dtype)
    bins = 0.5 * (z_ref[..., 1:] + z_ref[..., :-1])
    tf_shim.queue_uniform(sc['u_pdf'])
    s_ref = nu.sample_pdf(bins, w, S)
    s, above, below = O.sample_pdf(bins, w, S, sc['u_pdf'], return_indices=True)
    assert float((s - s_ref).abs().max()) < tol * 10
    # the reference's 63-step scan counts u >= cdf[k]: same integer bins
    cdf = torch.cumsum((w + 1e-5) / torch.sum(w + 1e-5, -1, keepdim=True), -1)
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], -1)
    above_def = (sc['u_pdf'][..., None] >= cdf[..., None, :]).sum(-1)
    if dtype == torch.float64:
        assert torch.equal(above.to(torch.int64), above_def)


def test_layers_match_reference_source(ref):
    from oracle.ref_runner import build_reference_nets
    _, ly, mv = ref
    p = {k: v.double() for k, v in O.init_params(2, bias_std=0.05).items()}
    nets = build_reference_nets(ly, p, 3)
    g = torch.Generator().manual_seed(0)
    pts = torch.randn(6, 4, 8, 3, generator=g).double()
    dirs = torch.randn(6, 4, 8, 3, generator=g).double()
    feats = torch.randn(6, 4, 8, 259, generator=g).double()
    emb, ro = nets['fine']
    emb.complete_output = True
    outs_ref = emb((pts, dirs, feats))
    outs = O.mv_embedding(pts, dirs, feats, p, 'fine_embedding', 3, complete_output=True)
    assert len(outs_ref) == len(outs) == 8
    for a, b in zip(outs, outs_ref):
        assert a.shape == b.shape and float((a - b).abs().max()) < 1e-11
    c_ref, s_ref = ro(outs_ref[-1])
    c, s = O.render_readout(outs[-1], p, 'fine_readout')
    assert float((c - c_ref).abs().max()) < 1e-12 and float((s - s_ref).abs().max()) < 1e-12
    rgb_ref, dep_ref, w_ref = mv.MVVNeRFRenderer.volumetric_render(torch.sort(torch.rand(2, 3, 8, generator=g).double()).values,
                                                                    s_ref[:2, :3], c_ref[:2, :3])
    z = torch.sort(torch.rand(2, 3, 8, generator=torch.Generator().manual_seed(0)).double()).values  # noqa: F841


@pytest.mark.parametrize('nv', [1, 3])
def test_call_matches_reference_source_fp64(ref, nv):
    """the whole _call, outputs AND gradients (reference autodiff == torch autograd through the stand-in)"""
    from oracle.ref_runner import reference_call
    sc = O.synthetic_scene(33, 1, nv, 6, 20, 28, dtype=torch.float64)
    base = O.init_params(4, bias_std=0.05, dtype=torch.float64)
    results = []
    for fn in ('ref', 'oracle'):
        p = {k: v.clone().requires_grad_(True) for k, v in base.items()}
        feats = sc['features'].clone().requires_grad_(True)
        if fn == 'ref':
            out = reference_call(sc['inputs'], feats, p, near=NEAR, far=FAR, n_samples=64, u_strat=sc['u_strat'],
                                 u_pdf=sc['u_pdf'])
        else:
            out = O.mvnerf_call(sc['inputs'], feats, p, near=NEAR, far=FAR, n_samples=64, u_strat=sc['u_strat'],
                                u_pdf=sc['u_pdf'])
        O.train_loss(out, sc['target']).backward()
        results.append((out, {k: v.grad for k, v in p.items()}, feats.grad))
    (o1, g1, f1), (o2, g2, f2) = results
    for a, b in zip(o1, o2):
        assert a.shape == b.shape and float((a - b).abs().max()) < 1e-10
    for k in g1:
        assert float((g1[k] - g2[k]).abs().max()) <= 1e-8 * max(1.0, float(g1[k].abs().max())), k
    assert float((f1 - f2).abs().max()) <= 1e-8 * max(1.0, float(f1.abs().max()))


# ---------------------------------------------------------------- pose-query models (delta_ngf / grasp_mvnerf)
def test_transforms_to_check_and_matrices_identical():
    from oracle import ref_runner as R
    p = {**O.init_params(0), **O.init_grasp_readout(1)}
    model = R.build_reference_query_model(p, 1, 1, 4)
    assert torch.equal(model.transforms_to_check[0, :, 0], O.transforms_to_check())
    for rep in ('quaternion', '6d'):
        t, r = O.synthetic_poses(0, 2, 5, rep)
        assert torch.equal(R.reference_compute_matrices(t, r, rep), O.compute_matrices(t, r, rep))


@pytest.mark.parametrize('which', ['delta_ngf', 'grasp_mvnerf'])
@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-10), (torch.float32, 5e-5)])
def test_query_call_matches_reference_source(which, dtype, tol):
    """DeltaNGF._call (delta_ngf/model.py:188-245) / GraspMVNeRF._call (grasp_mvnerf/model.py:113-158) and the
    gradient of the summed grasp success w.r.t. the pose matrices, from the reference's own source"""
    from oracle import ref_runner as R
    sc = O.synthetic_scene(8, 2, 3, 6, 20, 28, dtype=dtype)
    p = {**O.init_params(3, bias_std=0.02, dtype=dtype), **O.init_grasp_readout(4, bias_std=0.02, dtype=dtype)}
    _, _, img, K, E = sc['inputs']
    t, r = O.synthetic_poses(2, 2, 5, dtype=dtype)
    Mo = O.compute_matrices(t, r).requires_grad_(True)
    Mr = Mo.detach().clone().requires_grad_(True)
    yo = O.query_call(img, K, E, Mo, sc['features'], p)
    yr = R.reference_query_call(img, K, E, Mr, sc['features'], p, which)
    assert yo.shape == yr.shape == (2, 5)
    yo.sum().backward()
    yr.sum().backward()
    scale = float(yr.abs().max())
    assert float((yo - yr).abs().max()) / scale < tol
    assert float((Mo.grad - Mr.grad).abs().max() / Mr.grad.abs().max()) < (1e-8 if dtype == torch.float64 else 2e-2)
