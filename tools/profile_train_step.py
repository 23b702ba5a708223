#!/usr/bin/env python
"""torch.profiler breakdown (CUDA kernel time by name) of DeltaNGF.train_step at the reference's training size"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from clipnerf_b200 import synthetic  # noqa: E402
from src.lib.delta_ngf.model import DeltaNGF  # noqa: E402

dev = torch.device('cuda:0')
H, W, C = 480, 640, 256
b, nv, n_pts = 1, 3, 192
sc = synthetic.scene(1, 2, nv, H, W, C, device=dev)
_, _, img, K, E = [t.to(dev) for t in sc['inputs']]
feats = sc['features']
model = DeltaNGF(n_views=nv, batch_size=b, n_points_train=n_pts, device=dev)
model.compile()
model.feature_fn = lambda flat: feats.reshape(b * nv, H, W, C)
g = torch.Generator().manual_seed(0)
ctr = torch.tensor([0.6, 0.0, 0.1])


def poses():
    q = torch.randn(b, n_pts, 4, generator=g)
    return ctr + (torch.rand(b, n_pts, 3, generator=g) - 0.5) * 0.1, q / q.norm(dim=-1, keepdim=True)


lp, tp = poses(), poses()
data = ((lp[0], lp[1], tp[0], tp[1], img, K, E),
        (torch.rand(b, n_pts, generator=g), torch.randn(b, n_pts, 3, generator=g), torch.randn(b, n_pts, 4, generator=g)))
for _ in range(3):
    model.train_step(data)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        model.train_step(data)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=45, max_name_column_width=70))
