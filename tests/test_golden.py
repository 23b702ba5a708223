"""Golden vectors produced by the reference's own source (oracle/make_golden.py, build container):
the oracle must reproduce them on any machine (CPU), and so must the CUDA path (GPU, bf16-operand
tolerance)."""
import glob
import os

import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import make_golden as G
from oracle import nerf_oracle as O


def _cases():
    return sorted(os.path.basename(f)[:-4] for f in glob.glob(os.path.join(G.GOLDEN_DIR, '*.npz')))


def test_golden_files_present():
    assert _cases() == sorted(list(G.CASES) + list(G.QUERY_CASES))


@pytest.mark.parametrize('case', sorted(G.CASES))
def test_oracle_reproduces_reference_golden(case):
    gold = np.load(os.path.join(G.GOLDEN_DIR, f'{case}.npz'))
    inputs, feats, u_strat, u_pdf, target, params = G.make_inputs(case)
    p = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    fe = feats.clone().requires_grad_(True)
    out, aux = O.mvnerf_call(inputs, fe, p, near=G.NEAR, far=G.FAR, n_samples=64, u_strat=u_strat, u_pdf=u_pdf,
                             return_aux=True)
    loss = O.train_loss(out, target)
    loss.backward()
    assert np.array_equal(aux['z'].numpy(), gold['z_coarse'])                      # stratified z: bit-exact
    for k, v in zip(('rgb', 'depth', 'fine_rgb', 'fine_depth'), out):
        assert rel_err(v, gold[k]) < 2e-5, k
    assert abs(float(loss) - float(gold['loss'])) < 1e-6
    # gradients: fp32 on both sides, different summation order inside matmul / cumsum.  The coarse net
    # (and the feature map) receive the ill-conditioned term that flows through the sampler
    # (DESIGN.md section 2): fp32 rounding noise alone moves it by a few per cent, while in fp64 the
    # oracle and the reference source agree to 1e-8 (test_oracle_vs_reference_source.py).
    for k in p:
        tol = 5e-3 if k.startswith('fine') else 8e-2
        if 'grad.' + k in gold:
            assert rel_err(p[k].grad, gold['grad.' + k]) < tol, k
        else:
            assert abs(float(p[k].grad.norm()) / float(gold['gradl2.' + k]) - 1) < tol, k
            assert rel_err(p[k].grad.flatten()[::257], gold['gradsample.' + k]) < tol, k
    assert abs(float(fe.grad.norm()) / float(gold['grad_features_l2']) - 1) < 8e-2


@pytest.mark.gpu
@pytest.mark.parametrize('case', sorted(G.CASES))
def test_cuda_path_against_reference_golden(cuda, case):
    from clipnerf_b200 import params as P
    from clipnerf_b200 import render
    gold = np.load(os.path.join(G.GOLDEN_DIR, f'{case}.npz'))
    inputs, feats, u_strat, u_pdf, target, params = G.make_inputs(case)
    pc = P.from_named(params, 'coarse').to(cuda)
    pf = P.from_named(params, 'fine').to(cuda)
    out = render.render_rays(pc, pf, [t.to(cuda) for t in inputs], feats.to(cuda), u_strat.to(cuda), u_pdf.to(cuda),
                             G.NEAR, G.FAR)
    # golden = the reference source in PLAIN fp32.  Coarse outputs: 1e-2.  The fine pass of an 8-ray, single-view
    # case on a 12x16 white-noise map also inherits moved fine sample positions (the coarse weights that drive the
    # inverse CDF are computed with bf16 operands), measured 1.1e-2 on nv1_r8: 1.5e-2 here, and 1e-2 against the
    # oracle that rounds operands where the kernels do (same z_fine to ~1e-6).
    ref = O.mvnerf_call(inputs, feats, params, near=G.NEAR, far=G.FAR, n_samples=64, u_strat=u_strat, u_pdf=u_pdf,
                        quant='bf16')
    for k, v, r in zip(('rgb', 'depth', 'fine_rgb', 'fine_depth'), out, ref):
        e_gold, e_bf16 = rel_err(v, gold[k]), rel_err(v, r)
        print(f'[golden {case}] {k:10s} vs fp32 golden {e_gold:.2e}  vs bf16-operand oracle {e_bf16:.2e}')
        assert e_gold < (1.5e-2 if k.startswith('fine') else 1e-2), (k, e_gold)
        assert e_bf16 < 1e-2, (k, e_bf16)


@pytest.mark.parametrize('qcase', sorted(G.QUERY_CASES))
def test_oracle_reproduces_reference_query_golden(qcase):
    """DeltaNGF._call + d(sum grasp success)/d(pose variables), produced by the reference's own source"""
    gold = np.load(os.path.join(G.GOLDEN_DIR, f'{qcase}.npz'))
    img, K, E, feats, tr, rot, params = G.make_query_inputs(qcase)
    tr, rot = tr.clone().requires_grad_(True), rot.clone().requires_grad_(True)
    M = O.compute_matrices(tr, rot, G.QUERY_CASES[qcase]['rep'])
    assert rel_err(M, gold['matrices']) < 1e-6
    y = O.query_call(img, K, E, M, feats, params)
    y.sum().backward()
    assert rel_err(y, gold['grasp_success']) < 5e-5
    # fp32 on both sides with different summation orders; the pose gradient runs through posenc frequencies up to
    # pi * 2^9 rad/m, so fp32 rounding alone moves it by ~1e-2 (in fp64 oracle and reference source agree to 1e-8,
    # test_oracle_vs_reference_source.py::test_query_call_matches_reference_source)
    assert rel_err(tr.grad, gold['grad_translations']) < 2e-2
    assert rel_err(rot.grad, gold['grad_rotations']) < 2e-2


@pytest.mark.gpu
@pytest.mark.parametrize('qcase', sorted(G.QUERY_CASES))
def test_cuda_query_against_reference_golden(cuda, qcase):
    from clipnerf_b200 import params as P
    from src.lib.delta_ngf.model import DeltaNGF
    gold = np.load(os.path.join(G.GOLDEN_DIR, f'{qcase}.npz'))
    img, K, E, feats, tr, rot, params = G.make_query_inputs(qcase)
    c = G.QUERY_CASES[qcase]
    model = DeltaNGF(n_views=img.shape[1], batch_size=1, n_points_train=c['n_pts'], rotation_representation=c['rep'],
                     device=cuda)
    with torch.no_grad():
        model.fine_embedding.flat.copy_(P.from_named(params, 'fine')[:model.fine_embedding.flat.numel()].to(cuda))
    model.grasp_readout.load_keras_variables(params)
    model.set_pose(tr, rot)
    y = model._call((None,) * 4 + (img, K, E), model.compute_matrices(), c['n_pts'], feats.to(cuda))
    assert rel_err(y, gold['grasp_success']) < 1e-2, rel_err(y, gold['grasp_success'])
