"""N4: the torch mirrors of the reference's feature producers (``src/lib/mvnerf/layers.py:7-259, 414-660``).
CPU tests of the Keras semantics the mirrors write out, of the output shapes the hot path expects, of the
pre-upsample mode against the oracle's upsample, and of the Keras-layout variable exchange."""
import numpy as np
import torch

from src.lib.mvnerf import encoders as E


def test_same_padding_is_tensorflows():
    # TF 'same': out = ceil(n / s), the odd pad goes to the END (layers.py:47-48: 7x7 stride 2 on 480 x 640)
    x = torch.arange(480 * 640, dtype=torch.float32).view(1, 1, 480, 640)
    p = E.same_pad(x, 7, 2)
    assert p.shape[-2:] == (485, 645)
    assert float(p[0, 0, 2, 2]) == 0.0 and float(p[0, 0, 1, 2]) == 0.0 and float(p[0, 0, 2, 1]) == 0.0   # 2 before ...
    assert float(p[0, 0, 481, 641]) == float(x[0, 0, 479, 639]) and float(p[0, 0, 484, 644]) == 0.0      # ... 3 after
    conv = E.Conv2D(1, 1, 7, strides=2, padding='same', use_bias=False)
    assert conv(x).shape[-2:] == (240, 320)
    y = E.Conv2D(1, 1, 3, strides=2, padding='same')(torch.zeros(1, 1, 14, 14))         # post_process_4: 14 -> 7
    assert y.shape[-2:] == (7, 7)
    assert E.same_pad(torch.zeros(1, 1, 14, 14), 3, 2).shape[-2:] == (15, 15)            # one row / column, at the end


def test_bilinear_resize_is_the_half_pixel_upsample_of_the_oracle():
    from oracle import nerf_oracle as O
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 6, 8, 5, generator=g)                                  # NHWC
    up = E._nhwc(E.resize_bilinear(E._nchw(x), (12, 16)))
    assert torch.allclose(up, O.upsample2x_bilinear(x), atol=1e-6)


def test_multi_head_attention_matches_the_explicit_form():
    g = torch.Generator().manual_seed(1)
    m = E.MultiHeadAttention(32, 4, 8)
    x = torch.randn(2, 7, 32, generator=g)
    kv = E.keras_variables(m)
    q = torch.einsum('bte,ehd->bthd', x, kv['query.kernel']) + kv['query.bias']
    k = torch.einsum('bte,ehd->bthd', x, kv['key.kernel']) + kv['key.bias']
    v = torch.einsum('bte,ehd->bthd', x, kv['value.kernel']) + kv['value.bias']
    a = torch.softmax(torch.einsum('bthd,bshd->bhts', q, k) / np.sqrt(8.0), dim=-1)
    ref = torch.einsum('bhts,bshd,hde->bte', a, v, kv['output.kernel']) + kv['output.bias']
    assert torch.allclose(m(x), ref, atol=1e-5)


def test_block_shares_one_batchnorm_and_always_uses_batch_statistics():
    blk = E.Block(4, 4).eval()
    assert sum(1 for _ in blk.modules() if isinstance(_, E.BatchNorm)) == 1
    x = torch.randn(3, 4, 6, 6)
    y = blk(x)
    h = blk.conv_1(x)
    hn = (h - h.mean((0, 2, 3), keepdim=True)) / torch.sqrt(h.var((0, 2, 3), unbiased=False, keepdim=True) + 1e-3)
    h2 = blk.conv_2(torch.relu(hn))
    h2n = (h2 - h2.mean((0, 2, 3), keepdim=True)) / torch.sqrt(h2.var((0, 2, 3), unbiased=False, keepdim=True) + 1e-3)
    assert torch.allclose(y, torch.relu(h2n + x), atol=1e-5)


def test_transformer_block_adds_the_block_input_twice():
    blk = E.TransformerBlock(2, 16, 2).eval()
    x = torch.randn(2, 5, 16)
    a = x + blk.attention(blk.layer_norm_1(x.transpose(1, 2)).transpose(1, 2))
    assert torch.allclose(blk(x), x + blk.mlp(blk.layer_norm_2(a)), atol=1e-6)


def test_visual_features_shape_and_variable_exchange():
    torch.manual_seed(0)
    vf = E.VisualFeatures(256, (480, 640)).eval()
    assert hasattr(vf, 'vision_transformer') and hasattr(vf, 'conv_features')         # train_nerf.py:30-31 optimizer groups
    with torch.no_grad():
        out = vf(torch.rand(1, 480, 640, 3))
    assert out.shape == (1, 240, 320, 256) and torch.isfinite(out).all()
    kv = E.keras_variables(vf)
    assert kv['conv_features.conv_features.0.kernel'].shape == (7, 7, 3, 64)          # HWIO
    assert kv['vision_transformer.vit.transformer_blocks.0.0.attention.query.kernel'].shape == (768, 12, 64)
    assert kv['vision_transformer.post_process_1.1.kernel'].shape == (4, 4, 48, 48)   # Conv2DTranspose HWOI
    assert kv['vision_transformer.vit.pos_embedding'].shape == (1, 197, 768)
    # a checkpoint written from these variables restores another instance exactly
    saved = {k: v.detach().clone().numpy() for k, v in kv.items()}
    other = E.VisualFeatures(256, (480, 640)).eval()
    assert E.load_keras_variables(other, saved) == []
    for (n, a), (_, b_) in zip(sorted(E.keras_variables(vf).items()), sorted(E.keras_variables(other).items())):
        assert torch.equal(a, b_), n


def test_combine_clip_visual_v4_and_its_pre_upsample_form():
    from oracle import nerf_oracle as O
    torch.manual_seed(1)
    full = E.CombineCLIPVisualV4().eval()
    low = E.CombineCLIPVisualV4(final_upsample=False).eval()
    low.load_state_dict(full.state_dict())
    g = torch.Generator().manual_seed(2)
    clip = (torch.randn(1, 1024, generator=g), torch.randn(1, 56, 56, 256, generator=g), torch.randn(1, 28, 28, 512, generator=g),
            torch.randn(1, 14, 14, 1024, generator=g), torch.randn(1, 7, 7, 2048, generator=g))
    vis, text = torch.randn(1, 240, 320, 256, generator=g), torch.ones(1, 1024)
    with torch.no_grad():
        a, b_ = full((clip, vis, text)), low((clip, vis, text))
    assert a.shape == (1, 480, 640, 256) and b_.shape == (1, 240, 320, 256)
    # the reference's last layer is exactly the upsample the gather's low-resolution mode composes (layers.py:659)
    assert torch.allclose(a, O.upsample2x_bilinear(b_), atol=1e-5)
    assert E.CombineCLIPVisualV3().up_3.double_conv.conv_2.out_channels == 256 and full.up_3.double_conv.conv_2.out_channels == 128
    kv = E.keras_variables(full)
    assert kv['conv.kernel'].shape == (3, 3, 2048, 1024) and kv['conv_fusion_3.conv.kernel'].shape == (1, 1, 384, 256)
    assert not any('dense' in k for k in kv)                                       # use_dense=False: Slice, no variables
    assert E.keras_variables(E.CombineCLIPVisualV4(use_dense=True))['multiply_fusion_2.dense.kernel'].shape == (1024, 512)


def test_converter_name_resolver_and_npz_roundtrip(tmp_path):
    """tools/convert_tf_checkpoint.py walks the reference's Keras objects by the mirrors' variable names (its resolver is
    exercised on a stand-in object tree: TensorFlow is absent here); the ``.npz`` it writes is what
    ``MVVNeRFRenderer.load_reference_npz`` reads."""
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location('conv', os.path.join(root, 'tools', 'convert_tf_checkpoint.py'))
    conv = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(conv)

    class Obj:
        pass
    r = Obj()
    r.blocks = [Obj()]
    r.blocks[0].layers = [Obj()]
    blk = r.blocks[0].layers[0]
    blk.mlp, blk.attention = Obj(), Obj()
    blk.mlp.layers = ['dense_0', 'dense_1']
    blk.attention._output_dense = 'out'
    r.multiply_fusion_2 = Obj()
    r.multiply_fusion_2.tile = Obj()
    r.multiply_fusion_2.tile.dense = 'tile_dense'
    assert conv.resolve(r, 'blocks.0.0.mlp.2') == 'dense_1' and conv.resolve(r, 'blocks.0.0.mlp.0') == 'dense_0'
    assert conv.resolve(r, 'blocks.0.0.attention.output') == 'out'
    assert conv.resolve(r, 'multiply_fusion_2.dense') == 'tile_dense'

    from src.lib.mvnerf.model_v4 import MVVNeRFRenderer
    torch.manual_seed(0)
    a = MVVNeRFRenderer(64, 64, n_views=1)
    names = a.checkpoint_names()
    assert len(names) == len(set(names)) and 'combine_clip_visual.up_3.double_conv.conv_1.kernel' in names
    saved = {}
    for g in ('visual_features', 'combine_clip_visual'):
        saved.update({k: v.detach().numpy() for k, v in E.keras_variables(getattr(a, g), g + '.').items()})
    with torch.no_grad():
        a.nerf_params.normal_()
    for net in ('coarse', 'fine'):
        for sub in ('embedding', 'readout'):
            for i, (_, v) in enumerate(getattr(a, f'{net}_{sub}').variables().items()):
                saved[f'{net}_{sub}/{i:03d}'] = v.detach().numpy()
    np.savez(tmp_path / 'ref.npz', **saved)
    b_ = MVVNeRFRenderer(64, 64, n_views=1)
    assert b_.load_reference_npz(tmp_path / 'ref.npz')
    assert torch.equal(a.nerf_params, b_.nerf_params)
    for (n, x), (_, y) in zip(sorted(E.keras_variables(a.visual_features).items()), sorted(E.keras_variables(b_.visual_features).items())):
        assert torch.equal(x, y), n
