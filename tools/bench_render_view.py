#!/usr/bin/env python
"""cfg4 of BASELINE.json: full-image render (forward only + uint8 bookkeeping) through `render_view`; one JSON line.
    python tools/bench_render_view.py [--size 224x224|480x640] [--nv 1]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]
import numpy as np  # noqa: E402
import torch  # noqa: E402

from clipnerf_b200 import _abi, synthetic  # noqa: E402
if os.environ.get('CN_LIB'):            # another build of the library (before / after comparisons)
    _abi.LIB_PATH = os.path.abspath(os.environ['CN_LIB'])
from src.lib.mvnerf.model_v0 import MVVNeRFRenderer, render_view  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--size', default='224x224')
ap.add_argument('--nv', type=int, default=1)
ap.add_argument('--steps', type=int, default=5)
args = ap.parse_args()
H, W = [int(x) for x in args.size.split('x')]
dev = torch.device('cuda:0')
rs = np.random.RandomState(0)
poses = [synthetic.polar_pose(rs.uniform(0, 6.28), rs.uniform(0.2, 1.1), rs.uniform(0.7, 0.9)) for _ in range(args.nv + 1)]
K3 = synthetic.intrinsics(H, W)
cams = [dict(pose=p, intrinsics=np.asarray(K3).reshape(-1)) for p in poses]
colors = [(rs.uniform(0, 255, (H, W, 3))).astype(np.uint8) for _ in range(args.nv)]
model = MVVNeRFRenderer(512, 512, n_views=args.nv, near=0.3, far=1.3, device=dev)
feats = torch.randn(args.nv, H, W, 256, device=dev)
model.feature_fn = lambda images: feats
for _ in range(2):
    render_view(model, colors, cams[:-1], cams[-1])
torch.cuda.synchronize()
l0 = _abi.LAUNCHES
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(args.steps):
    rgb8, depth8 = render_view(model, colors, cams[:-1], cams[-1])
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / args.steps
print(json.dumps({'case': 'render_view', 'metric': 'rays_per_sec_fwd', 'value': H * W / (ms * 1e-3), 'unit': 'rays/s',
                  'ms_per_image': ms, 'gpu_launches_per_image': (_abi.LAUNCHES - l0) // args.steps,
                  'd2h_bytes_per_image': int(rgb8.nbytes + depth8.nbytes),
                  'config': {'workload': f'cfg4: render_view {H}x{W} ({H * W} rays x (64+128) samples), nv={args.nv}, forward only, '
                                         f'device ray generation + uint8 conversion, includes host image upload and uint8 download'}}))
