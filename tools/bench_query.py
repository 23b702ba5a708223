#!/usr/bin/env python
"""Timing of the pose-query path; prints one JSON line per case.

    python tools/bench_query.py [--steps K] [--warmup W]

  * optimize_pose : DNGFOptimizer.optimize_pose at the reference's validation size -- 3 source images (nv = 1,
                    480x640x256 fp32 maps), 4096 pose guesses x 42 query offsets = 516 096 query rows per step,
                    forward + pose-gradient backward + optimizer step (grasp_optimizer.py:158-184)
  * train_step    : DeltaNGF.train_step at the reference's training size -- b = 1, nv = 3, np = 192 poses
                    (8 064 query points, 24 192 per-view rows), two forwards, pose gradient, tangent pass (forward kernels with the primal masks), head update
Device-timed with CUDA events; synthetic scene, random-init weights."""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]

import torch  # noqa: E402

H, W, C = 480, 640, 256


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--guesses', type=int, default=4096)
    ap.add_argument('--case', default='all', choices=['all', 'optimize_pose', 'train_step'])
    args = ap.parse_args()
    from clipnerf_b200 import _abi, synthetic
    from src.lib.delta_ngf.grasp_optimizer import DNGFOptimizer
    from src.lib.delta_ngf.model import DeltaNGF
    import torch.distributed as dist
    rank, world = int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        # under torch.distributed.run: every rank trains on its own scene / poses, DeltaNGF.train_step averages the
        # GraspReadout gradient over the ranks (one all-reduce of 2.33 MB); device time = max over ranks
        os.environ.setdefault('NCCL_DEBUG_FILE', '/dev/stderr')
        dist.init_process_group('nccl', device_id=dev)
        args.case = 'train_step'
    _abi.load()
    torch.manual_seed(0)          # identical initial weights on every rank (a DDP wrapper would broadcast them)
    bounds = [[0.55, 0.65], [-0.05, 0.05], [0.05, 0.15]]

    if args.case in ('all', 'optimize_pose'):
        n_img, n_guess = 3, args.guesses
        sc = synthetic.scene(0, 2, n_img, H, W, C, device=dev)           # [1, 3, ...]: 3 images, regrouped to b=3, nv=1
        _, _, img, K, E = [t.to(dev) for t in sc['inputs']]
        model = DeltaNGF(n_views=1, batch_size=1, n_points_train=n_guess, device=dev)
        opt = DNGFOptimizer(model, bounds, n_initial_guesses=n_guess, n_images=n_img, clip_translation=True)
        opt.compile()
        opt.set_initial_guesses(opt.generate_initial_guesses())
        inputs, feats = [img, K, E, None], sc['features']
        l0 = _abi.LAUNCHES
        ms = timed(lambda: opt.optimize_pose(inputs, feats, (True, True)), args.steps, args.warmup)
        launches = (_abi.LAUNCHES - l0) // (args.steps + args.warmup)
        rows = n_img * n_guess * 42
        replay = opt.capture(inputs, feats, (True, True))
        ms_g = timed(replay, args.steps, args.warmup)
        print(json.dumps({'case': 'optimize_pose (CUDA graph replay)', 'metric': 'query_points_per_sec_fwd_bwd',
                          'value': rows / (ms_g * 1e-3), 'unit': 'points/s', 'ms_per_step': ms_g, 'rows_per_step': rows,
                          'graph': 'DNGFOptimizer.capture(): one step recorded with torch.cuda.graph, replayed'}))
        print(json.dumps({'case': 'optimize_pose', 'metric': 'query_points_per_sec_fwd_bwd', 'value': rows / (ms * 1e-3),
                          'unit': 'points/s', 'ms_per_step': ms, 'rows_per_step': rows, 'gpu_launches_per_step': launches,
                          'mlp_tflops': rows * 491264.0 * 3 / (ms * 1e-3) / 1e12,
                          'config': {'workload': f'DNGFOptimizer.optimize_pose: {n_img} images x {n_guess} guesses x 42 '
                                                 f'offsets, nv=1, {H}x{W}x{C} fp32 maps, fwd + pose gradient + Adam'}}))

    if args.case in ('all', 'train_step'):
        b, nv, n_pts = 1, 3, 192
        sc = synthetic.scene(1 + rank, 2, nv, H, W, C, device=dev)
        _, _, img, K, E = [t.to(dev) for t in sc['inputs']]
        feats = sc['features']
        model = DeltaNGF(n_views=nv, batch_size=b, n_points_train=n_pts, device=dev)
        model.compile()
        model.feature_fn = lambda flat: feats.reshape(b * nv, H, W, C)
        g = torch.Generator().manual_seed(rank)
        ctr = torch.tensor([0.6, 0.0, 0.1])

        def poses():
            q = torch.randn(b, n_pts, 4, generator=g)
            return ctr + (torch.rand(b, n_pts, 3, generator=g) - 0.5) * 0.1, q / q.norm(dim=-1, keepdim=True)
        lp, tp = poses(), poses()
        data = ((lp[0], lp[1], tp[0], tp[1], img, K, E),
                (torch.rand(b, n_pts, generator=g), torch.randn(b, n_pts, 3, generator=g), torch.randn(b, n_pts, 4, generator=g)))
        l0 = _abi.LAUNCHES
        if world > 1:
            dist.barrier()
        ms = timed(lambda: model.train_step(data), args.steps, args.warmup)
        launches = (_abi.LAUNCHES - l0) // (args.steps + args.warmup)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
            w0 = [p.detach().clone() for p in model.grasp_readout.parameters()]
            for p in w0:                                  # every rank holds the same head after the averaged steps
                ref = p.clone()
                dist.broadcast(ref, 0)
                assert torch.equal(ref, p), 'ranks diverged'
        if world == 1:
            replay = model.capture_train_step(data)
            ms_g = timed(lambda: replay(data), args.steps, args.warmup)
            print(json.dumps({'case': 'train_step (CUDA graph replay)', 'metric': 'poses_per_sec_train_step',
                              'value': b * n_pts / (ms_g * 1e-3), 'unit': 'poses/s', 'ms_per_step': ms_g,
                              'graph': 'DeltaNGF.capture_train_step(): the batch is copied into static buffers, one graph launch per step'}))
        if rank == 0:
            print(json.dumps({'case': 'train_step', 'metric': 'poses_per_sec_train_step',
                          'value': world * b * n_pts / (ms * 1e-3), 'n_gpus': world,
                          'unit': 'poses/s', 'ms_per_step': ms, 'query_points_per_step': 2 * world * b * n_pts * 42,
                          'gpu_launches_per_step': launches,
                          'config': {'workload': f'DeltaNGF.train_step: b={b}, nv={nv}, np={n_pts} (x42 offsets), landscape + '
                                                 f'gradient-matching losses, tangent pass in torch, Adam on GraspReadout'}}))

    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
