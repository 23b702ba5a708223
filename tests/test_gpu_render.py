"""End-to-end parity of the CUDA render path with the oracle's ``MVVNeRFRenderer._call``:
forward outputs (rel-err <= 1e-2) and gradients of MSE(coarse)+MSE(fine) w.r.t. every NeRF
parameter and the feature maps (rel-err <= 2e-2)."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu

NEAR, FAR = 0.3, 1.3


def _run(cuda, sc, named, need_feat_grad=True):
    from clipnerf_b200 import params, render
    pc = params.from_named(named, 'coarse').to(cuda).requires_grad_(True)
    pf = params.from_named(named, 'fine').to(cuda).requires_grad_(True)
    inputs = [t.to(cuda) for t in sc['inputs']]
    feats = sc['features'].to(cuda).requires_grad_(need_feat_grad)
    out = render.render_rays(pc, pf, inputs, feats, sc['u_strat'].to(cuda), sc['u_pdf'].to(cuda), NEAR, FAR)
    return out, pc, pf, feats


@pytest.mark.parametrize('nv', [1, 3])
def test_render_forward_backward(cuda, nv):
    from clipnerf_b200 import params
    from oracle import nerf_oracle as O
    sc = O.synthetic_scene(11, 1, nv, 32, 48, 64)
    named = O.init_params(5, bias_std=0.02)
    # oracle in fp32 (the reference's dtype), gradients by autograd
    p = {k: v.clone().requires_grad_(True) for k, v in named.items()}
    feats = sc['features'].clone().requires_grad_(True)
    ref, aux = O.mvnerf_call(sc['inputs'], feats, p, near=NEAR, far=FAR, n_samples=64, u_strat=sc['u_strat'],
                             u_pdf=sc['u_pdf'], return_aux=True)
    O.train_loss(ref, sc['target']).backward()

    out, pc, pf, fd = _run(cuda, sc, named)
    names = ('rgb', 'depth', 'fine_rgb', 'fine_depth')
    for n_, a, b_ in zip(names, out, ref):
        e = rel_err(a, b_.detach())
        assert e < 1e-2, f'{n_}: rel err {e}'
    tgt = sc['target'].to(cuda)
    loss = torch.mean((out[0] - tgt) ** 2) + torch.mean((out[2] - tgt) ** 2)
    loss.backward()
    for net, flat in (('coarse', pc), ('fine', pf)):
        for name, gv in params.to_named(flat.grad.cpu(), net).items():
            e = rel_err(gv, p[name].grad)
            assert e < 2e-2, f'{name}: grad rel err {e}'
    e = rel_err(fd.grad, feats.grad)
    assert e < 2e-2, f'feature-map grad rel err {e}'


def test_render_inference_matches_training_forward(cuda):
    from oracle import nerf_oracle as O
    sc = O.synthetic_scene(12, 2, 2, 16, 32, 40)
    named = O.init_params(6, bias_std=0.02)
    with torch.no_grad():
        out_inf, *_ = _run(cuda, sc, named, need_feat_grad=False)
    out_trn, *_ = _run(cuda, sc, named)
    for a, b_ in zip(out_inf, out_trn):
        assert torch.equal(a, b_.detach())
