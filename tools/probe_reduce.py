"""Throughput of adding 81 KiB tiles to a [3,480,640,256] fp32 map: cp.reduce.async.bulk (mode 0)
vs red.global.add.v4.f32 (mode 1).  Run on a B200: python tools/probe_reduce.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'thesis-clip-nerf_b200'))
from clipnerf_b200 import ops  # noqa: E402

dev = torch.device('cuda:0')
m = torch.zeros(3, 480, 640, 256, device=dev)
for mode in (0, 1):
    for cps in (1, 2):
        iters = 100
        ops.probe_tile_reduce(m, 10, mode, cps)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.probe_tile_reduce(m, iters, mode, cps)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        tiles = 148 * cps * iters
        print(f'mode {mode} ctas/sm {cps}: {ms:.3f} ms for {tiles} tiles = {tiles * 82944 / ms / 1e6:.1f} GB/s of tile payload, '
              f'{ms * 1e3 / iters:.2f} us per tile per CTA')
print('sum check', float(m.sum()))
