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


def test_host_adam_follows_keras_and_the_reference_optimizer_groups():
    """RenderTrainer.apply_gradients (torch ops, host scalars) against a literal restatement of optimize()
    (nerf_utils.py:8-12) + tf.keras Adam + WarmupScheduler (nerf_utils.py:288-300) in float64; readouts are in no
    optimizer group of the reference (train_nerf.py:27-33) and keep their values unless train_readouts=True"""
    import numpy as np
    from clipnerf_b200._abi import EMB_PARAMS
    from clipnerf_b200.trainer import RenderTrainer
    dev = torch.device('cpu')
    for train_readouts in (False, True):
        tr = RenderTrainer(dev, seed=2, warmup_steps=3, target_lr=1e-3, train_readouts=train_readouts)
        p = tr.params.detach().double().numpy().copy()
        m, v = np.zeros_like(p), np.zeros_like(p)
        g = torch.Generator().manual_seed(0)
        for t in range(1, 7):
            grad = torch.randn(tr.params.shape, generator=g) * (3.0 if t == 3 else 0.1)      # some entries beyond the clip
            tr.params.grad = grad.clone()
            tr.apply_gradients()
            gc = np.clip(grad.double().numpy(), -1.0, 1.0)
            lr = O.warmup_lr(t, 1e-3, 3, 450000)
            m = 0.9 * m + 0.1 * gc
            v = 0.999 * v + 0.001 * gc * gc
            upd = lr * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t) * m / (np.sqrt(v) + 1e-7)
            if not train_readouts:
                upd[:, EMB_PARAMS:] = 0.0
            p = p - upd
            assert tr.iterations == t
            err = np.abs(tr.params.detach().double().numpy() - p).max() / np.abs(p).max()
            assert err < 1e-6, (t, err)
        if not train_readouts:
            init = RenderTrainer(dev, seed=2).params.detach()
            assert torch.equal(tr.params.detach()[:, EMB_PARAMS:], init[:, EMB_PARAMS:])
