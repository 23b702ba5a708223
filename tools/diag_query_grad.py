"""diagnostic: distribution of the per-point error of d/d(points) of the query path against the oracle"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]
import torch
from clipnerf_b200 import params, query
from oracle import nerf_oracle as O

cuda = torch.device('cuda:0')
b, nv, n = 1, 3, 1024
sc = O.synthetic_scene(3, b, nv, 8, 48, 64)
named = O.init_params(7, bias_std=0.02)
_, _, img, K, E = sc['inputs']
g = torch.Generator().manual_seed(5)
pts = torch.tensor([0.6, 0.0, 0.1]) + (torch.rand((b, n, 3), generator=g) - 0.5) * 0.12
d = torch.randn((b, n, 3), generator=g); dirs = d / d.norm(dim=-1, keepdim=True)
for taps in ([0], [3], [0, 1, 2, 3]):
    gw = [torch.randn(b, n, 128, generator=torch.Generator().manual_seed(20 + i)) * (1.0 if i in taps else 0.0) for i in range(4)]
    res = {}
    for quant in (None, 'bf16'):
        po, do = pts.clone().requires_grad_(True), dirs.clone().requires_grad_(True)
        out = O.query_embeddings_from_points(img, K, E, po, do, sc['features'], named, quant=quant)
        sum((o * w).sum() for o, w in zip(out, gw)).backward()
        res[quant] = (po.grad, do.grad)
    pf = params.from_named(named, 'fine').to(cuda)
    pc, dc = pts.to(cuda).requires_grad_(True), dirs.to(cuda).requires_grad_(True)
    out = query.query_embeddings(pf, img.to(cuda), K.to(cuda), E.to(cuda), sc['features'].to(cuda), pc, dc)
    sum((o * w.to(cuda)).sum() for o, w in zip(out, gw)).backward()
    for name, got, idx in (('points', pc.grad.cpu(), 0), ('dirs', dc.grad.cpu(), 1)):
        for q in (None, 'bf16'):
            ref = res[q][idx]
            e = (got - ref).norm(dim=-1)[0] / ref.norm(dim=-1)[0].max()
            l2 = float((got - ref).norm() / ref.norm())
            qs = torch.quantile(e, torch.tensor([0.5, 0.9, 0.99]))
            print(f'taps {taps} grad {name:6s} vs oracle[{q}]: L2 {l2:.3e}  per-row/max med {qs[0]:.2e} p90 {qs[1]:.2e} p99 {qs[2]:.2e} max {e.max():.2e}')
    e32 = res['bf16'][0] - res[None][0]
    print(f'   oracle bf16 vs fp32: L2 {float(e32.norm() / res[None][0].norm()):.3e}')

# ---- module level: per-pose gradient of DeltaNGF
from src.lib.delta_ngf.model import DeltaNGF
b, nv, n_pts = 2, 3, 64
sc = O.synthetic_scene(4, b, nv, 8, 48, 64)
named = O.init_params(9, bias_std=0.02)
named.update(O.init_grasp_readout(2, bias_std=0.02))
_, _, img, K, E = sc['inputs']
for rep in ('quaternion', '6d'):
    t, r = O.synthetic_poses(1, b, n_pts, rep)
    res = {}
    for q in (None, 'bf16'):
        to, ro = t.clone().requires_grad_(True), r.clone().requires_grad_(True)
        ref = O.query_call(img, K, E, O.compute_matrices(to, ro, rep), sc['features'], named, quant=q)
        ref.sum().backward()
        res[q] = (to.grad, ro.grad, ref.detach())
    model = DeltaNGF(n_views=nv, batch_size=b, n_points_train=n_pts, rotation_representation=rep, device=cuda)
    with torch.no_grad():
        model.fine_embedding.flat.copy_(params.from_named(named, 'fine')[:model.fine_embedding.flat.numel()].to(cuda))
    model.grasp_readout.load_keras_variables(named)
    model.set_pose(t, r)
    y = model._call((None,) * 4 + (img, K, E), model.compute_matrices(), n_pts, sc['features'].to(cuda))
    gt, gr = torch.autograd.grad(y.sum(), model.pose_variables)
    for name, got, idx in (('t', gt.cpu(), 0), ('r', gr.cpu(), 1)):
        for q in (None, 'bf16'):
            ref = res[q][idx]
            e = (got - ref).norm(dim=-1).reshape(-1) / ref.norm(dim=-1).max()
            qs = torch.quantile(e, torch.tensor([0.5, 0.9]))
            print(f'{rep} grad {name} vs oracle[{q}]: L2 {float((got - ref).norm() / ref.norm()):.3e} per-pose med {qs[0]:.2e} p90 {qs[1]:.2e} max {e.max():.2e}')
    e = res['bf16'][0] - res[None][0]
    print(f'   oracle bf16 vs fp32 grad t: L2 {float(e.norm() / res[None][0].norm()):.3e}; y err {float((y.cpu() - res["bf16"][2]).abs().max() / res["bf16"][2].abs().max()):.2e}')
