#!/usr/bin/env python
"""device time of the tcgen05 GraspReadout kernels alone at the optimize_pose size (3 x 4096 poses x 42 offsets)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]
import torch
from clipnerf_b200 import _abi, ops

dev = torch.device('cuda:0')
b, n_poses = 3, int(sys.argv[1]) if len(sys.argv) > 1 else 4096
n = n_poses * 42
n_pad = (n + 127) // 128 * 128
taps = torch.randn(4, b * n_pad, 128, device=dev) * 0.7
pack = ops.grasp_head_pack(torch.randn(_abi.HEAD_PARAMS, device=dev) * 0.05)
gy = torch.randn(b, n_poses, device=dev)
for elu in (True, False):
    records = []
    import contextlib

    @contextlib.contextmanager
    def timer(name):
        a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); yield; c.record(); records.append((name, a, c))
    for it in range(6):
        if it == 3:
            _abi.TIMER = timer
        y, st = ops.grasp_head_fwd(taps, b, n_pad, n_poses, pack, elu, True)
        d = ops.grasp_head_bwd(gy, st)
    torch.cuda.synchronize()
    _abi.TIMER = None
    acc = {}
    for name, a, c in records:
        acc[name] = acc.get(name, 0.0) + a.elapsed_time(c) / 3
    print('elu' if elu else 'relu', {k: round(v, 3) for k, v in acc.items()}, 'ms; rows', b * n_pad)
