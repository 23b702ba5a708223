"""model_v4.MVVNeRFRenderer on the GPU: the torch feature producers feeding the sm_100a render path, full-resolution
maps against the pre-upsample maps + folded gather (N4 producer of the N2 mode)."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


def test_model_v4_call_with_and_without_the_final_upsample(cuda):
    from oracle import nerf_oracle as O
    from src.lib.mvnerf.model_v4 import MVVNeRFRenderer
    sc = O.synthetic_scene(3, 1, 2, 64, 480, 640, C=8)                 # rays / cameras / images at the real image size
    g = torch.Generator().manual_seed(5)
    pyramid = [torch.randn(2, 1024, generator=g), torch.randn(2, 56, 56, 256, generator=g), torch.randn(2, 28, 28, 512, generator=g),
               torch.randn(2, 14, 14, 1024, generator=g), torch.randn(2, 7, 7, 2048, generator=g)]
    clip = lambda images: tuple(t.to(images.device) for t in pyramid)  # noqa: E731  (the CLIP SavedModel is a plug-in)
    torch.manual_seed(0)
    full = MVVNeRFRenderer(64, 64, n_views=2, near=0.3, far=1.3, clip_visual=clip, device=cuda)
    low = MVVNeRFRenderer(64, 64, n_views=2, near=0.3, far=1.3, clip_visual=clip, lowres_features=True, device=cuda)
    low.load_state_dict(full.state_dict())
    for m in (full, low):
        m.eval()
        m._generator = torch.Generator(device=cuda).manual_seed(11)    # same stratified / inverse-CDF draws
    with torch.no_grad():
        f_full = full.encode(sc['inputs'][2].reshape(2, 480, 640, 3))
        f_low = low.encode(sc['inputs'][2].reshape(2, 480, 640, 3))
        assert f_full.shape == (2, 480, 640, 256) and f_low.shape == (2, 240, 320, 256)
        a = full.call(sc['inputs'])
        b_ = low.call(sc['inputs'])
    for name, x, y in zip(('rgb', 'depth', 'fine_rgb', 'fine_depth'), a, b_):
        assert x.shape == y.shape and torch.isfinite(x).all()
        print(f'[model_v4] {name}: folded vs full-resolution {rel_err(y, x):.2e}')
        assert rel_err(y, x) < 1e-2


def test_language_nerf_call_runs_the_reference_chain(cuda):
    """LanguageNeRF(encoders=True): CLIP plug-ins -> VisualFeatures -> CombineCLIPVisualV4(use_dense, elu) -> the sm_100a
    pose-query path (lmvnerf/model_v4.py:178-187), full-resolution maps against the pre-upsample maps."""
    from oracle import nerf_oracle as O
    from src.lib.lmvnerf.model_v4 import LanguageNeRF
    b, nv, n_pts = 1, 2, 6
    sc = O.synthetic_scene(3, b, nv, 8, 480, 640, C=8)
    g = torch.Generator().manual_seed(5)
    pyramid = [torch.randn(nv, 1024, generator=g), torch.randn(nv, 56, 56, 256, generator=g), torch.randn(nv, 28, 28, 512, generator=g),
               torch.randn(nv, 14, 14, 1024, generator=g), torch.randn(nv, 7, 7, 2048, generator=g)]
    text = torch.randn(nv, 1024, generator=g)
    clip_v = lambda images: tuple(t.to(images.device) for t in pyramid)   # noqa: E731
    clip_t = lambda tokens: text.to(cuda)                                 # noqa: E731
    torch.manual_seed(0)
    models = [LanguageNeRF(n_views=nv, batch_size=b, n_points_train=n_pts, encoders=True, clip_visual=clip_v, clip_textual=clip_t,
                           lowres_features=low, device=cuda).eval() for low in (False, True)]
    models[1].load_state_dict(models[0].state_dict())
    t, r = O.synthetic_poses(0, b, n_pts)
    _, _, img, K, E = sc['inputs']
    outs = []
    for m in models:
        m.set_pose(t, r)
        with torch.no_grad():
            outs.append(m.call((None, None, None, None, img, K, E, torch.zeros(nv, 77, dtype=torch.long))))
    assert outs[0].shape == (b, n_pts) and torch.isfinite(outs[0]).all()
    print('[LanguageNeRF] folded vs full-resolution maps', rel_err(outs[1], outs[0]))
    assert rel_err(outs[1], outs[0]) < 1e-2
