"""Host-side logic that needs no GPU: flat parameter layout, tile-image addressing."""
import torch

from oracle import nerf_oracle as O


def test_flat_parameter_layout_roundtrip():
    from clipnerf_b200 import _abi, params
    named = O.init_params(3, bias_std=0.1)
    for net in ('coarse', 'fine'):
        flat = params.from_named(named, net)
        assert flat.numel() == _abi.NET_PARAMS == 247300
        back = params.to_named(flat, net)
        assert set(back) == {k for k in named if k.startswith(net)}
        for k, v in back.items():
            assert torch.equal(v, named[k])
    tab = params.variable_table()
    assert tab[0] == ('embedding.layer_0.kernel', (379, 128), 0)
    assert tab[-1][0] == 'readout.output_layer.bias' and tab[-1][2] == 247296
    # same order / shapes as the oracle's (Keras) variable list
    names = [n for n, _, _ in tab]
    assert names[:26] == ['embedding.' + n for n, _ in O.embedding_param_shapes()]


def test_tile_image_roundtrip_and_swizzle():
    from clipnerf_b200 import layout
    m = torch.randn(256, 384).to(torch.bfloat16).to(torch.float32)
    img = layout.matrix_to_tiles(m)
    assert img.numel() == 2 * 98304
    assert torch.equal(layout.tiles_to_matrix(img, 384), m)
    # row r, 16-byte chunk c sits at chunk position c ^ (r & 7) of its 128-byte row
    r, c = 13, 3
    off = (c // 8) * 16384 + r * 128 + (((c % 8) ^ (r % 8)) * 16)
    got = img[off:off + 16].view(torch.bfloat16).to(torch.float32)
    assert torch.equal(got, m[r, c * 8:c * 8 + 8])


def test_x_column_order_is_a_permutation():
    from clipnerf_b200 import layout
    cols = [layout.x_col_to_ref(j) for j in range(384)]
    assert sorted(cols[:379]) == list(range(379)) and cols[379:] == [-1] * 5
    assert cols[0] == 123 and cols[256] == 0 and cols[316] == 60 and cols[376] == 120


def test_device_side_adam_matches_host_side():
    """RenderTrainer.apply_gradients_on_device (step counter, warm-up schedule and bias corrections as device scalars:
    the form a CUDA graph can replay) against apply_gradients, same gradients, across the warm-up boundary"""
    import torch
    from clipnerf_b200.trainer import RenderTrainer
    dev = torch.device('cpu')
    a, b = RenderTrainer(dev, seed=2, warmup_steps=3), RenderTrainer(dev, seed=2, warmup_steps=3)
    g = torch.Generator().manual_seed(0)
    for i in range(6):
        grad = torch.randn(a.params.shape, generator=g) * (3.0 if i == 2 else 0.1)      # some entries beyond the clip
        a.params.grad, b.params.grad = grad.clone(), grad.clone()
        a.apply_gradients()
        b.apply_gradients_on_device()
        assert float(b._t_dev) == a.iterations == i + 1
        err = float((a.params - b.params).abs().max() / a.params.abs().max())
        assert err < 1e-6, (i, err)
