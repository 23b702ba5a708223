"""The reference's module surface on the GPU: op-level functions of nerf_utils.py with their own
signatures, and MVVNeRFRenderer / render_view as train_nerf.py uses them."""
import numpy as np
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu

NEAR, FAR = 0.3, 1.3


def test_op_level_functions_match_oracle(cuda):
    from oracle import nerf_oracle as O
    from src.lib.mvnerf import nerf_utils as U
    from src.lib.mvnerf.model_v0 import MVVNeRFRenderer
    sc = O.synthetic_scene(21, 2, 3, 16, 24, 32)
    ro, rd, img, K, E = sc['inputs']
    d = lambda t: t.to(cuda)  # noqa: E731
    pts_r, z_r = O.sample_along_ray(ro, rd, NEAR, FAR, 64, sc['u_strat'])
    pts, z = U.sample_along_ray(d(ro), d(rd), NEAR, FAR, 2, 16, 64, u=d(sc['u_strat']))
    assert torch.equal(z.cpu(), z_r) and rel_err(pts, pts_r) < 1e-6
    pix_r, cam_r = O.compute_pixel_in_image_mv(pts_r, K, E)
    pix, cam = U.compute_pixel_in_image_mv(d(pts_r), d(K), d(E))
    assert torch.equal(pix.cpu(), pix_r) and torch.equal(cam.cpu(), cam_r)
    dirs_r = O.world_to_camera_direction_vector_mv(rd, E)
    dirs = U.world_to_camera_direction_vector_mv(d(rd), d(E), 3)
    assert torch.equal(dirs.cpu(), dirs_r)
    g_r = O.get_projection_features_mv(img * 2 - 1.0, sc['features'], pix_r)
    g = U.get_projection_features_mv(d(img * 2 - 1.0), d(sc['features']), d(pix_r), 16, 64, 2)
    assert torch.equal(g.cpu(), g_r)
    enc_r = O.position_encoding(cam_r[..., :3].reshape(6, 16, 64, 3))
    enc = U.position_encoding(d(cam_r[..., :3].reshape(6, 16, 64, 3)), 10, np.pi)
    assert enc.shape == enc_r.shape and float((enc.cpu() - enc_r).abs().max()) < 2e-4
    w = torch.rand(2, 16, 62, generator=torch.Generator().manual_seed(1))
    bins = 0.5 * (z_r[..., 1:] + z_r[..., :-1])
    s_r = O.sample_pdf(bins, w, 64, sc['u_pdf'])
    s = U.sample_pdf(d(bins), d(w), 64, u=d(sc['u_pdf']))
    assert torch.equal(s.cpu(), s_r)
    sig = torch.rand(2, 16, 64) * 10
    col = torch.rand(2, 16, 64, 3)
    rgb_r, dep_r, w_r = O.volumetric_render(z_r, sig, col)
    rgb, dep, ww = MVVNeRFRenderer.volumetric_render(d(z_r), d(sig), d(col))
    assert rel_err(rgb, rgb_r) < 1e-5 and rel_err(dep, dep_r) < 1e-5 and rel_err(ww, w_r) < 1e-5
    a = U.sigma_to_alpha(d(sig), d(z_r))
    assert rel_err(a, O.sigma_to_alpha(sig, z_r)) < 1e-6


def _model(cuda, nv, named=None):
    from clipnerf_b200 import params
    from src.lib.mvnerf.model_v0 import MVVNeRFRenderer
    m = MVVNeRFRenderer(n_rays_train=32, n_rays_infer=32, n_views=nv, near=NEAR, far=FAR,
                        original_image_size=(24, 32), device=cuda)
    if named is not None:
        with torch.no_grad():
            for net in ('coarse', 'fine'):
                flat = params.from_named(named, net).to(cuda)
                getattr(m, f'{net}_embedding').flat.copy_(flat[:246784])
                getattr(m, f'{net}_readout').flat.copy_(flat[246784:])
    return m


def test_model_call_infer_and_train_step(cuda, tmp_path):
    from oracle import nerf_oracle as O
    sc = O.synthetic_scene(22, 1, 2, 32, 24, 32)
    named = O.init_params(3, bias_std=0.02)
    m = _model(cuda, 2, named)
    feats = sc['features'].to(cuda)
    ref = O.mvnerf_call(sc['inputs'], sc['features'], named, near=NEAR, far=FAR, n_samples=64, u_strat=sc['u_strat'],
                        u_pdf=sc['u_pdf'])
    inputs_np = tuple(t.numpy() for t in sc['inputs'])          # the generator hands numpy arrays
    out = m.infer(inputs_np, feats, sc['u_strat'], sc['u_pdf'])
    for a, b in zip(out, ref):
        assert rel_err(a, b) < 1e-2
    # call() goes through the plug-in feature producer
    m.feature_fn = lambda images: feats.reshape(-1, *feats.shape[2:])
    m.compile(warmup_steps=1)                                    # default optimizer: RenderTrainer on nerf_params
    from clipnerf_b200.trainer import RenderTrainer
    assert isinstance(m.optimizer, RenderTrainer) and m.optimizer.params is m.nerf_params
    before, ro_before = m.coarse_embedding.flat.detach().clone(), m.fine_readout.flat.detach().clone()
    losses = [float(m.train_step((inputs_np, sc['target'].numpy()))['loss']) for _ in range(8)]
    assert np.isfinite(losses).all() and losses[-1] < losses[0]
    assert not torch.equal(before, m.coarse_embedding.flat.detach())
    # the reference's optimizer groups hold the embeddings only (train_nerf.py:27-33): readouts keep their values
    assert torch.equal(ro_before, m.fine_readout.flat.detach()) and m.optimizer.iterations == 8
    # store / load round trip, False when a group is missing
    path = str(tmp_path / 'model_final')
    assert m.load(path) is False
    m.store(path)
    m2 = _model(cuda, 2)
    assert m2.load(path) is True
    assert torch.equal(m2.fine_embedding.flat, m.fine_embedding.flat)
    assert torch.equal(m2.coarse_readout.flat, m.coarse_readout.flat)
    # fit over a generator
    from clipnerf_b200.synthetic import SyntheticNerfDataset
    from src.lib.data_generator.mvnerf import MVNeRFDataGenerator
    ds = SyntheticNerfDataset(n_scenes=2, n_perspectives=4, H=24, W=32)
    gen = MVNeRFDataGenerator(ds, n_rays_train=32, batch_size=1, n_views=2, shuffle=True)
    m.feature_fn = lambda images: torch.tanh(images.mean(-1, keepdim=True)).expand(-1, -1, -1, 256).contiguous()
    hist = m.fit(gen, epochs=2, initial_epoch=0)
    assert len(hist['loss']) == 2 and np.isfinite(hist['loss']).all()


def test_fit_is_one_graph_launch_per_step(cuda, monkeypatch):
    """MVVNeRFRenderer.fit with the default optimizer (what src/train_nerf.py:52 calls) drives RenderTrainer: after the
    first (eager) step and one capture per device buffer set, every step is ONE CUDA-graph launch behind the single
    uint8 -> float conversion launch of the uploaded source images (the generator's compact batches); an explicit
    torch optimizer selects the plain eager step."""
    from clipnerf_b200 import _abi
    from clipnerf_b200.synthetic import SyntheticNerfDataset
    from src.lib.data_generator.mvnerf import MVNeRFDataGenerator
    ds = SyntheticNerfDataset(n_scenes=6, n_perspectives=4, H=24, W=32)
    np.random.seed(3)
    gen = MVNeRFDataGenerator(ds, n_rays_train=32, batch_size=1, n_views=2, shuffle=True)
    m = _model(cuda, 2)
    maps = torch.nn.Parameter(torch.randn(2, 24, 32, 256, device=cuda))          # "encoder output" that trains itself
    m.feature_fn = lambda images: maps
    m.compile(warmup_steps=2)
    replays = []
    real_replay = torch.cuda.CUDAGraph.replay
    monkeypatch.setattr(torch.cuda.CUDAGraph, 'replay', lambda self: (replays.append(_abi.LAUNCHES), real_replay(self))[1])
    p0 = m.nerf_params.detach().clone()
    hist = m.fit(gen, epochs=1)                      # 6 steps: 1 eager, 2 captures (+ replay), 3 pure replays
    assert len(replays) == 5 and m.optimizer.iterations == 6 and len(m.loss_tracker) == 6
    # between two replays the library launches exactly one kernel outside the graph: cn_images_u8_to_f32
    assert replays[3] - replays[2] == 1 and replays[4] - replays[3] == 1, replays
    launches_end = _abi.LAUNCHES
    assert np.isfinite(hist['loss']).all() and not torch.equal(p0, m.nerf_params.detach())
    assert maps.grad is not None and torch.isfinite(maps.grad).all() and float(maps.grad.abs().max()) > 0
    assert float(m.optimizer._t_dev) == 6.0
    hist = m.fit(gen, epochs=3, initial_epoch=1)     # graphs are reused across epochs
    assert len(replays) == 5 + 12 and _abi.LAUNCHES == launches_end + 12 and m.optimizer.iterations == 18
    # explicit torch optimizer: eager autograd step over the same flat parameter
    m2 = _model(cuda, 2)
    m2.feature_fn = lambda images: maps.detach()
    m2.compile(optimizer=torch.optim.Adam(m2.parameters(), lr=1e-3, eps=1e-7))
    assert m2._trainer is None
    q0 = m2.nerf_params.detach().clone()
    m2.fit(gen, epochs=1)
    assert not torch.equal(q0, m2.nerf_params.detach())


def test_render_view_bookkeeping(cuda):
    """full-image render: pixel order, odd-size chunking and the truncating uint8 conversion are exact
    given the float outputs; the floats agree with the oracle within 1e-2."""
    from clipnerf_b200.synthetic import SyntheticNerfDataset
    from oracle import nerf_oracle as O
    from src.lib.mvnerf.model_v0 import render_view
    H, W = 15, 21                                   # 315 rays: odd, and not a multiple of the chunk
    ds = SyntheticNerfDataset(n_scenes=1, n_perspectives=3, H=H, W=W, seed=4)
    colors = [ds.datasets['color'].read_sample_at_idx(0, i) for i in range(2)]
    cams = [ds.datasets['camera_config'].read_sample_at_idx(0, i) for i in range(2)]
    tgt = ds.datasets['camera_config'].read_sample_at_idx(0, 2)
    named = O.init_params(4, bias_std=0.02)
    m = _model(cuda, 2, named)
    g = torch.Generator().manual_seed(0)
    feats = torch.randn(1, 2, H, W, 256, generator=g)
    m.feature_fn = lambda images: feats.to(cuda).reshape(-1, H, W, 256)
    u_s, u_p = torch.rand(1, H * W + 1, 64, generator=g), torch.rand(1, H * W + 1, 64, generator=g)
    # fixed draws: patch the model's RNG so that both chunks see the rows of (u_s, u_p) they cover
    calls = []

    def fake_rand(b, nr):
        calls.append(nr)
        k = len(calls) - 1
        src = u_s if k % 2 == 0 else u_p
        start = sum(calls[:k][::2]) if k % 2 == 0 else sum(calls[:k][1::2])
        return src[:, start:start + nr].to(cuda)

    m._rand = fake_rand
    rgb8, depth8 = render_view(m, colors, cams, tgt, chunk=200)
    assert rgb8.shape == (H, W, 3) and rgb8.dtype == np.uint8 and depth8.shape == (H, W, 1)
    # oracle on the same rays (row-major, u = column)
    K3 = np.reshape(tgt['intrinsics'], (3, 3)).astype(np.float32)
    o, d = O.get_rays(W, H, tgt['pose'], K3)
    inputs = O.get_input([c[..., :3] for c in colors], cams, d.reshape(-1, 3), o.reshape(-1, 3))
    # chunk 1 = rays [0,200), chunk 2 = rays [200,315) padded by one copy of the last ray
    us = torch.cat([u_s[:, :200], u_s[:, 200:315]], 1)
    up = torch.cat([u_p[:, :200], u_p[:, 200:315]], 1)
    ref = O.mvnerf_call([torch.tensor(a) for a in inputs], feats, named, near=NEAR, far=FAR, n_samples=64,
                        u_strat=us, u_pdf=up)
    ref8, refd8 = O.quantize_render(ref[2][0].numpy(), ref[3][0].numpy(), (H, W))
    diff = np.abs(rgb8.astype(int) - ref8.astype(int))
    assert diff.max() <= 4 and (diff <= 1).mean() > 0.9, (diff.max(), (diff <= 1).mean())
    assert np.abs(depth8.astype(int) - refd8.astype(int)).max() <= 6


def test_device_rays_and_uint8_conversion(cuda):
    """cn_generate_rays against the reference's numpy get_rays (pixel order exact, values to an fp32 ulp) and
    cn_quantize_render against the reference's uint8 bookkeeping (bit-exact)."""
    from clipnerf_b200 import ops
    from oracle import nerf_oracle as O
    H, W = 37, 53
    pose = O.polar_pose(0.9, 0.6, 0.8)                                   # float64, like the dataset's camera configs
    K = np.array([[55.4, 0, W / 2], [0, 55.4, H / 2], [0, 0, 1]], dtype=np.float32)
    ro, rd = O.get_rays(W, H, pose, K)
    for p in (pose, pose.astype(np.float32)):
        ro_ref, rd_ref = O.get_rays(W, H, p, K)
        o, d = ops.generate_rays(p, K, W, 0, H * W, cuda)
        assert np.array_equal(o.cpu().numpy(), ro_ref.reshape(-1, 3).astype(np.float32))
        err = np.abs(d.cpu().numpy() - rd_ref.reshape(-1, 3).astype(np.float32)).max()
        assert err <= (1.2e-7 if p.dtype == np.float64 else 5e-7), err
    # a pixel range in the middle of the image starts at the right pixel
    o, d = ops.generate_rays(pose, K, W, 1000, 64, cuda)
    assert np.abs(d.cpu().numpy() - rd.reshape(-1, 3)[1000:1064].astype(np.float32)).max() <= 1.2e-7
    g = torch.Generator().manual_seed(1)
    rgb = torch.rand(H * W, 3, generator=g) * 1.2 - 0.1                   # values below 0 and above 1
    depth = torch.rand(H * W, generator=g) * 0.9 + 0.3
    rgb8, depth8 = ops.quantize_render(rgb.to(cuda), depth.to(cuda))
    ref8, refd8 = O.quantize_render(rgb.numpy(), depth.numpy(), (H, W))
    assert np.array_equal(rgb8.cpu().numpy().reshape(H, W, 3), ref8)
    assert np.array_equal(depth8.cpu().numpy().reshape(H, W, 1), refd8)


def test_compact_batches_equal_the_reference_batches(cuda):
    """MVNeRFDataGenerator.compact_item (uint8 source images, converted by cn_images_u8_to_f32 on the device) against
    __getitem__ (the reference's host expression np.array(colors) / 255.0 -> float32): same numpy RNG draws, every
    array bit-identical"""
    from clipnerf_b200 import ops
    from clipnerf_b200.synthetic import SyntheticNerfDataset
    from src.lib.data_generator.mvnerf import MVNeRFDataGenerator
    ds = SyntheticNerfDataset(n_scenes=3, n_perspectives=4, H=30, W=41)
    gen = MVNeRFDataGenerator(ds, n_rays_train=64, batch_size=2, n_views=3, shuffle=False)
    np.random.seed(5)
    ref_in, ref_t = gen[0]
    np.random.seed(5)
    cmp_in, cmp_t = gen.compact_item(0)
    assert cmp_in[2].dtype == np.uint8 and cmp_in[2].shape == (2, 3, 30, 41, 4)
    for k in (0, 1, 3, 4):
        assert np.array_equal(ref_in[k], cmp_in[k])
    assert np.array_equal(ref_t, cmp_t)
    img = ops.images_u8_to_f32(torch.as_tensor(cmp_in[2]).to(cuda))
    assert img.shape == (2, 3, 30, 41, 3) and np.array_equal(img.cpu().numpy(), ref_in[2])
