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
    assert _cases() == sorted(G.CASES)


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
    for k, v in zip(('rgb', 'depth', 'fine_rgb', 'fine_depth'), out):
        assert rel_err(v, gold[k]) < 1e-2, (k, rel_err(v, gold[k]))
