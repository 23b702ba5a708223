"""The module surface src/train_nerf.py consumes from src/lib: names, signatures,
host-side bookkeeping.  CPU-only parts here; the CUDA parts are in test_gpu_interface.py."""
import inspect

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O


def test_import_paths_and_signatures():
    from src.lib.data_generator.mvnerf import MVNeRFDataGenerator
    from src.lib.mvnerf.model_v0 import MVVNeRFRenderer, render_view
    from src.lib.mvnerf.nerf_utils import WarmupScheduler, load_pretrained_weights  # noqa: F401
    sig = inspect.signature(MVVNeRFRenderer.__init__).parameters
    for k in ('n_rays_train', 'n_rays_infer', 'n_views', 'n_samples', 'n_features', 'embed_direction_vector',
              'batch_size', 'near', 'far', 'original_image_size'):
        assert k in sig
    assert sig['near'].default == 0.7 and sig['far'].default == 1.5      # ctor defaults differ from the YAML
    for m in ('call', 'infer', '_call', 'volumetric_render', 'train_step', 'compile', 'fit', 'summary', 'store',
              'load', 'encode'):
        assert hasattr(MVVNeRFRenderer, m)
    assert list(inspect.signature(render_view).parameters)[:4] == ['model', 'src_colors', 'src_camera_configs',
                                                                   'tgt_camera_config']
    assert list(inspect.signature(MVNeRFDataGenerator.get_input).parameters) == ['colors', 'camera_configs', 'r_d',
                                                                                 'r_o']


def test_model_attributes_for_train_nerf():
    """train_nerf.py:27-32 builds optimizer groups from these attributes."""
    from src.lib.mvnerf.model_v0 import MVVNeRFRenderer
    m = MVVNeRFRenderer(n_rays_train=512, n_rays_infer=512, n_views=1, near=0.3, far=1.3)
    for a in ('coarse_embedding', 'fine_embedding', 'coarse_readout', 'fine_readout', 'combine_clip_visual',
              'clip_visual'):
        assert hasattr(m, a)
    assert hasattr(m.visual_features, 'vision_transformer') and hasattr(m.visual_features, 'conv_features')
    assert m.n_rays_infer == 512
    names = set(m.coarse_embedding.variables())
    assert 'layer_0.kernel' in names and 'fusion_blocks.2.layer_1.bias' in names and len(names) == 26
    assert tuple(m.coarse_embedding.variables()['layer_0.kernel'].shape) == (379, 128)
    assert sum(p.numel() for p in m.parameters()) == 2 * 247300
    k = m.fine_embedding.variables()['feature_blocks.0.layer_0.kernel']
    lim = (6.0 / 256) ** 0.5
    assert float(k.abs().max()) <= lim and float(k.abs().max()) > 0.9 * lim        # glorot_uniform
    assert float(m.fine_embedding.variables()['layer_0.bias'].abs().max()) == 0.0  # zeros
    with pytest.raises(ValueError):
        MVVNeRFRenderer(512, 512, n_samples=32)


def test_host_ray_bookkeeping_matches_oracle_bit_for_bit():
    from src.lib.data_generator.util import camera_parameters
    from src.lib.mvnerf import nerf_utils as U
    pose = O.polar_pose(0.4, 0.8, 0.85)
    K = np.array([[55.4, 0, 32], [0, 55.4, 24], [0, 0, 1]], dtype=np.float32)
    o1, d1 = U.get_rays(64, 48, pose, K)
    o2, d2 = O.get_rays(64, 48, pose, K)
    assert np.array_equal(o1, o2) and np.array_equal(d1, d2)
    np.random.seed(7)
    r1 = U.bbox_biased_sample(512, np.array([0, 0, 48, 64]), 48, 64)
    r2 = O.bbox_biased_sample(512, np.array([0, 0, 48, 64]), 48, 64, rng=np.random.RandomState(7))
    assert np.array_equal(r1, r2) and r1.dtype == r2.dtype
    e1, k1 = camera_parameters(dict(pose=pose, intrinsics=K.reshape(-1)))
    e2, k2 = O.camera_parameters(dict(pose=pose, intrinsics=K.reshape(-1)))
    assert np.array_equal(e1, e2) and np.array_equal(k1, k2)


def test_data_generator_batches():
    from clipnerf_b200.synthetic import SyntheticNerfDataset
    from src.lib.data_generator.mvnerf import MVNeRFDataGenerator
    ds = SyntheticNerfDataset(n_scenes=3, n_perspectives=4, H=24, W=32)
    np.random.seed(0)
    gen = MVNeRFDataGenerator(ds, n_rays_train=64, batch_size=2, n_views=3, shuffle=True)
    assert len(gen) == 1
    inputs, targets = gen[0]
    assert [a.shape for a in inputs] == [(2, 64, 3), (2, 64, 3), (2, 3, 24, 32, 3), (2, 3, 4, 4), (2, 3, 4, 4)]
    assert all(a.dtype == np.float32 for a in inputs) and targets.shape == (2, 64, 3)
    assert inputs[2].min() >= 0 and inputs[2].max() <= 1 and np.allclose(np.linalg.norm(inputs[1], axis=-1), 1)
    assert np.allclose(inputs[3][0, 0, 3], [0, 0, 0, 1]) and np.allclose(inputs[4][0, 0, 3], [0, 0, 0, 1])


def test_warmup_scheduler_matches_oracle():
    from src.lib.mvnerf.nerf_utils import WarmupScheduler
    s = WarmupScheduler(1e-4, 10000, 450000)
    for step in (0, 1, 5000, 10000, 10001, 450000, 450001):
        assert s(step) == pytest.approx(O.warmup_lr(step, 1e-4, 10000, 450000), rel=1e-12, abs=1e-18)


def test_layers_error_behaviour():
    from src.lib.mvnerf.layers import ResNetMLPBlock
    with pytest.raises(ValueError, match='not supported'):
        ResNetMLPBlock(128, 128, activation='tanh')
    blk = ResNetMLPBlock(16, 8, transform_shortcut=True, activation='elu', input_size=32)
    assert tuple(blk(torch.randn(5, 32)).shape) == (5, 8)


def test_pose_query_module_surface():
    """what train_delta_ngf.py / train_language.py / train_goal.py consume"""
    from src.lib.delta_ngf.grasp_optimizer import DNGFOptimizer
    from src.lib.delta_ngf.layers import GraspReadout
    from src.lib.delta_ngf.model import DeltaNGF
    from src.lib.grasp_mvnerf.model import GraspMVNeRF
    from src.lib.lmvnerf.model_v4 import LanguageNeRF
    sig = inspect.signature(DeltaNGF.__init__).parameters
    for k in ('n_features', 'original_image_size', 'n_points_train', 'n_views', 'n_5d_poses', 'pretrained',
              'batch_size', 'rotation_representation', 'softmax_before_loss'):
        assert k in sig
    for cls in (DeltaNGF, LanguageNeRF, GraspMVNeRF):
        for m in ('call', 'infer', '_call', 'train_step', 'compile', 'fit', 'store', 'load', 'load_backbone'):
            assert hasattr(cls, m), (cls, m)
    assert list(inspect.signature(DeltaNGF.infer).parameters)[1:] == ['inputs', 'transforms', 'n_points_infer',
                                                                       'batched_features']
    assert list(inspect.signature(GraspMVNeRF.infer).parameters)[1:] == ['inputs', 'n_points_infer', 'batched_features']
    m = DeltaNGF(n_views=3, batch_size=2, n_points_train=5)
    assert m.n_transforms_to_check == 42 and torch.equal(m.transforms_to_check, O.transforms_to_check())
    assert sum(p.numel() for p in m.grasp_readout.parameters()) == 582337
    assert hasattr(m.visual_features, 'vision_transformer') and m.n_views == 3
    assert tuple(m.translations.shape) == (2, 5, 3) and tuple(m.quaternions.shape) == (2, 5, 4)
    assert tuple(DeltaNGF(rotation_representation='6d').six_d_rotations.shape) == (1, 5, 6)
    with pytest.raises(ValueError, match='Unknown rotation representation'):
        DeltaNGF(rotation_representation='euler')
    with pytest.raises(NotImplementedError):
        DeltaNGF(pretrained=False)
    # pose matrices and the affordance head agree with the oracle on CPU tensors
    t, r = O.synthetic_poses(0, 2, 5)
    m.set_pose(t, r)
    assert torch.allclose(m.compute_matrices(), O.compute_matrices(t, r), atol=1e-7)
    named = O.init_grasp_readout(2, bias_std=0.02)
    m.grasp_readout.load_keras_variables(named)
    emb = [torch.randn(2, 5, 42, 128) for _ in range(4)]
    assert torch.allclose(m.grasp_readout(emb), O.grasp_readout(emb, named), atol=1e-4)
    assert GraspReadout(use_bias=False).readout[2].output_layer.bias is None
    opt = DNGFOptimizer(m, [[0, 1], [0, 1], [0, 1]], n_initial_guesses=8, n_images=3)
    g = opt.generate_initial_guesses()
    assert g[0].shape == (1, 8, 3) and g[1].shape == (1, 8, 4)
    opt.set_initial_guesses(g)
    assert len(opt.get_results()) == 8 and opt.get_results()[0].shape == (4, 4)
    for a in ('compile', 'optimize_pose', 'compute_current_grasp_success', 'post_process'):
        assert hasattr(opt, a)
