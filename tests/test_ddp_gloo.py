"""world_size-2 test of the one collective on the path (gradient all-reduce of the flat NeRF
parameter buffer) with the gloo backend on CPU: after the step every rank holds identical
parameters, equal to a single-process step on the mean gradient."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _grad_for(rank, n):
    g = torch.Generator().manual_seed(100 + rank)
    return torch.randn(2, n, generator=g) * 3.0          # some entries beyond the +-1 clip


def _worker(rank, world, port, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from clipnerf_b200._abi import NET_PARAMS
    from clipnerf_b200.trainer import RenderTrainer
    t = RenderTrainer(torch.device('cpu'), seed=5, warmup_steps=1)
    assert t.world == world
    for i in range(3):
        g = _grad_for(rank, NET_PARAMS)
        t._synced = False                         # (loss_and_backward resets it)
        if i == 1:
            # the form used on the GPU: the sum is started inside the backward (render.GRAD_SYNC hook, asynchronous),
            # allreduce_gradients then only takes the mean
            work = t._sync_inside_backward(g)
            work.wait()
        t.params.grad = g
        t.allreduce_gradients()
        t.apply_gradients()
    out[rank] = t.params.detach().clone()
    # the pose-query trainers' collective: mean of a list of gradients, None entries as zeros
    from clipnerf_b200.trainer import allreduce_mean
    like = [torch.zeros(3, 4), torch.zeros(5), torch.zeros(2, 2)]
    grads = [torch.full((3, 4), float(rank + 1)), None if rank == 0 else torch.ones(5), torch.full((2, 2), 2.0 * rank)]
    red = allreduce_mean(grads, like)
    assert torch.equal(red[0], torch.full((3, 4), 1.5)) and torch.equal(red[1], torch.full((5,), 0.5))
    assert torch.equal(red[2], torch.full((2, 2), 1.0))
    dist.destroy_process_group()


def test_gradient_allreduce_world2():
    import conftest  # noqa: F401  (sys.path)
    from clipnerf_b200._abi import NET_PARAMS
    from clipnerf_b200.trainer import RenderTrainer
    mgr = mp.Manager()
    out = mgr.dict()
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert torch.equal(out[0], out[1])
    ref = RenderTrainer(torch.device('cpu'), seed=5, warmup_steps=1)
    for _ in range(3):
        ref.params.grad = (_grad_for(0, NET_PARAMS) + _grad_for(1, NET_PARAMS)) / 2
        ref.apply_gradients()
    assert torch.allclose(out[0], ref.params.detach(), atol=1e-7)
    assert ref.iterations == 3
