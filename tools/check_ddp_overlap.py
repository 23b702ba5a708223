#!/usr/bin/env python
"""N-rank check of the gradient all-reduce started inside the backward (render.GRAD_SYNC) against the plain
all-reduce behind it: same mean gradient on every rank, and the step time of both forms.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/check_ddp_overlap.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    from clipnerf_b200 import _abi, synthetic
    from clipnerf_b200.trainer import RenderTrainer
    _abi.load()
    sc = synthetic.scene(rank, 4096, 3, device=dev)
    inputs = [t.to(dev) for t in sc['inputs']]
    u_s, u_p, tgt = sc['u_strat'].to(dev), sc['u_pdf'].to(dev), sc['target'].to(dev)
    feats = sc['features'].requires_grad_(True)
    res = {}
    for overlap in (False, True):
        tr = RenderTrainer(dev, seed=3, bias_std=0.02)
        tr.overlap_allreduce = overlap

        def step():
            feats.grad = None
            tr.loss_and_backward(inputs, feats, tgt, u_s, u_p, 0.3, 1.3)
            tr.allreduce_gradients()
        for _ in range(3):
            step()
        assert tr._synced == overlap
        g = tr.params.grad.clone()
        gathered = [torch.empty_like(g) for _ in range(world)]
        dist.all_gather(gathered, g)
        for other in gathered:
            assert torch.equal(other, g), 'ranks disagree on the reduced gradient'
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / 20], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        res[overlap] = (g, float(ms))
    err = float((res[True][0] - res[False][0]).abs().max() / res[False][0].abs().max())
    if rank == 0:
        print(f'world {world}: all-reduce behind the backward {res[False][1]:.3f} ms/step, inside it {res[True][1]:.3f} ms/step, '
              f'gradient difference {err:.2e}')
    assert err < 1e-4, err            # atomics order of the local backward only
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
