"""Golden vectors for the render path, produced by the reference's OWN source files running on the
TensorFlow stand-in (oracle/tf_shim + oracle/ref_runner) in fp32.  TEST INFRASTRUCTURE ONLY.

    python -m oracle.make_golden            # (build container only: needs /root/reference)

Inputs are NOT stored: they are regenerated from numpy's legacy ``RandomState`` (bit-stable across
numpy versions) by ``make_inputs``; the .npz files under tests/golden/ hold the reference outputs,
intermediate sample positions / bins and gradients of MSE(coarse)+MSE(fine)."""
from __future__ import annotations

import math
import os

import numpy as np
import torch

from . import nerf_oracle as O

NEAR, FAR = 0.3, 1.3
CASES = {
    'nv1_r8': dict(seed=101, nv=1, nr=8, H=12, W=16),
    'nv3_r8': dict(seed=103, nv=3, nr=8, H=12, W=16),
}
GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden')


def make_inputs(case: str):
    """(inputs 5-tuple, features, u_strat, u_pdf, target, params) as fp32 torch tensors"""
    c = CASES[case]
    rs = np.random.RandomState(c['seed'])
    nv, nr, H, W = c['nv'], c['nr'], c['H'], c['W']
    f = 0.5 * W / math.tan(math.radians(30.0))
    K3 = np.array([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1]])
    poses = [O.polar_pose(rs.uniform(0, 2 * math.pi), rs.uniform(0.2, 1.1), rs.uniform(0.7, 0.9)) for _ in range(nv + 1)]
    rc = np.stack([rs.randint(0, H, nr), rs.randint(0, W, nr)], axis=1)
    o, d = O.get_specific_rays(rc[:, 1], rc[:, 0], poses[-1], K3.astype(np.float32))
    e_invs, ks = zip(*[O.camera_parameters(dict(pose=p, intrinsics=K3.reshape(-1))) for p in poses[:-1]])
    t = lambda a: torch.tensor(np.asarray(a, dtype=np.float32))  # noqa: E731
    inputs = (t(o)[None], t(d)[None], t(rs.uniform(0, 1, (1, nv, H, W, 3))), t(ks)[None], t(e_invs)[None])
    feats = t(rs.standard_normal((1, nv, H, W, 256)))
    u_strat, u_pdf = t(rs.uniform(0, 1, (1, nr, 64))), t(rs.uniform(0, 1, (1, nr, 64)))
    target = t(rs.uniform(0, 1, (1, nr, 3)))
    params = {}
    for net in ('coarse', 'fine'):
        for name, shape in O.embedding_param_shapes() + O.readout_param_shapes():
            grp = 'embedding' if not name.startswith('output_layer') else 'readout'
            if name.endswith('kernel'):
                lim = math.sqrt(6.0 / (shape[0] + shape[1]))
                params[f'{net}_{grp}.{name}'] = t(rs.uniform(-lim, lim, shape))
            else:
                params[f'{net}_{grp}.{name}'] = t(rs.standard_normal(shape) * 0.02)
    return inputs, feats, u_strat, u_pdf, target, params


def generate():
    from .ref_runner import reference_call
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    for case in CASES:
        inputs, feats, u_strat, u_pdf, target, params = make_inputs(case)
        p = {k: v.clone().requires_grad_(True) for k, v in params.items()}
        fe = feats.clone().requires_grad_(True)
        out = reference_call(inputs, fe, p, near=NEAR, far=FAR, n_samples=64, u_strat=u_strat, u_pdf=u_pdf)
        loss = O.train_loss(out, target)
        loss.backward()
        # intermediates recomputed with the reference's own op functions
        from . import tf_shim
        nu, _, _ = tf_shim.load_reference()
        tf_shim.queue_uniform(u_strat)
        _, z = nu.sample_along_ray(inputs[0], inputs[1], NEAR, FAR, 1, inputs[0].shape[1], 64)
        rec = dict(rgb=out[0], depth=out[1], fine_rgb=out[2], fine_depth=out[3], loss=loss, z_coarse=z,
                   grad_features_l2=fe.grad.norm(), grad_features_sample=fe.grad.flatten()[::997])
        for k, v in p.items():
            if v.numel() <= 1024:
                rec['grad.' + k] = v.grad
            else:
                rec['gradl2.' + k] = v.grad.norm()
                rec['gradsample.' + k] = v.grad.flatten()[::257]
        np.savez_compressed(os.path.join(GOLDEN_DIR, f'{case}.npz'),
                            **{k: v.detach().numpy() for k, v in rec.items()})
        print(case, 'loss', float(loss), 'rgb[0]', out[0][0, 0].tolist())


QUERY_CASES = {
    'query_nv3_p6': dict(case='nv3_r8', seed=211, n_pts=6, rep='quaternion'),
    'query_nv1_p6_6d': dict(case='nv1_r8', seed=213, n_pts=6, rep='6d'),
}


def make_query_inputs(qcase: str):
    """(images, K, E^-1, features, translations, rotations, params incl. grasp_readout) of a pose-query case:
    the scene of the render case plus poses and GraspReadout weights from numpy's legacy RandomState"""
    c = QUERY_CASES[qcase]
    inputs, feats, _, _, _, params = make_inputs(c['case'])
    rs = np.random.RandomState(c['seed'])
    t = lambda a: torch.tensor(np.asarray(a, dtype=np.float32))  # noqa: E731
    tr = t(np.array([0.6, 0.0, 0.1]) + rs.uniform(-0.04, 0.04, (1, c['n_pts'], 3)))
    if c['rep'] == 'quaternion':
        q = rs.standard_normal((1, c['n_pts'], 4))
        rot = t(q / np.linalg.norm(q, axis=-1, keepdims=True))
    else:
        rot = t(rs.standard_normal((1, c['n_pts'], 6)))
    for name, shape in O.grasp_readout_param_shapes():
        params[f'grasp_readout.{name}'] = t(rs.standard_normal(shape) * (math.sqrt(2.0 / shape[0]) if name.endswith('kernel') else 0.02))
    return inputs[2], inputs[3], inputs[4], feats, tr, rot, params


def generate_query():
    """DeltaNGF._call from the reference's source (ref_runner.reference_query_call) + d(sum)/d(pose variables)"""
    from .ref_runner import reference_compute_matrices, reference_query_call
    for qcase, c in QUERY_CASES.items():
        img, K, E, feats, tr, rot, params = make_query_inputs(qcase)
        tr, rot = tr.clone().requires_grad_(True), rot.clone().requires_grad_(True)
        M = reference_compute_matrices(tr, rot, c['rep'])
        y = reference_query_call(img, K, E, M, feats, params)
        y.sum().backward()
        np.savez_compressed(os.path.join(GOLDEN_DIR, f'{qcase}.npz'), grasp_success=y.detach().numpy(),
                            matrices=M.detach().numpy(), grad_translations=tr.grad.numpy(),
                            grad_rotations=rot.grad.numpy())
        print(qcase, 'grasp_success', y.detach()[0, :3].tolist())


if __name__ == '__main__':
    generate()
    generate_query()
