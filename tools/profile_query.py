#!/usr/bin/env python
"""torch.profiler breakdown (CUDA kernel time by name) of one DNGFOptimizer.optimize_pose step at the validation size"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from clipnerf_b200 import synthetic  # noqa: E402
from src.lib.delta_ngf.grasp_optimizer import DNGFOptimizer  # noqa: E402
from src.lib.delta_ngf.model import DeltaNGF  # noqa: E402

dev = torch.device('cuda:0')
n_img, n_guess = 3, int(sys.argv[1]) if len(sys.argv) > 1 else 4096
sc = synthetic.scene(0, 2, n_img, 480, 640, 256, device=dev)
_, _, img, K, E = sc['inputs']
model = DeltaNGF(n_views=1, batch_size=1, n_points_train=n_guess, device=dev)
opt = DNGFOptimizer(model, [[0.55, 0.65], [-0.05, 0.05], [0.05, 0.15]], n_initial_guesses=n_guess, n_images=n_img)
opt.compile()
opt.set_initial_guesses(opt.generate_initial_guesses())
inputs, feats = [img, K, E, None], sc['features']
for _ in range(3):
    opt.optimize_pose(inputs, feats, (True, True))
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        opt.optimize_pose(inputs, feats, (True, True))
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=30, max_name_column_width=70))
