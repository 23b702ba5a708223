"""Where does the CUDA-graph training loop leave the eager loop's trajectory?  

 (a) eager vs eager, same seeds: the run-to-run noise floor of losses / parameters
 (b) graph vs eager, graph vs graph
 (c) optimizer alone ON THE GPU: the same gradient tensor fed to apply_gradients (host scalars) and
     apply_gradients_on_device (device scalars), per step, across the warm-up boundary
 (d) the same, but the device-side update replayed from a CUDA graph
 (e) per-step gradient of the graph replay vs an eager backward at the SAME parameters

    python tools/diag_graph_loop.py [--lr 1e-4] [--steps 7]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')):
    sys.path.insert(0, p)

from clipnerf_b200.trainer import RenderTrainer, fit_host_batches, fit_host_batches_graph  # noqa: E402
from oracle import nerf_oracle as O  # noqa: E402

NEAR, FAR = 0.3, 1.3


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--lr', type=float, default=1e-4)
    ap.add_argument('--steps', type=int, default=7)
    ap.add_argument('--rays', type=int, default=32)
    args = ap.parse_args()
    dev = torch.device('cuda:0')
    batches = []
    for seed in (41, 42, 43):
        sc = O.synthetic_scene(seed, 1, 2, args.rays, 48, 64)
        batches.append([t.pin_memory() for t in list(sc['inputs']) + [sc['target'], sc['u_strat'], sc['u_pdf']]])
    feats = O.synthetic_scene(41, 1, 2, args.rays, 48, 64)['features'].to(dev)
    seq = [batches[i % 3] for i in range(args.steps)]

    def run(fit):
        tr = RenderTrainer(dev, seed=3, bias_std=0.02, warmup_steps=4, target_lr=args.lr)
        f = feats.clone().requires_grad_(True)
        losses = fit(tr, iter(seq), f, NEAR, FAR)
        return losses, tr.params.detach().clone()

    runs = {'eager0': run(fit_host_batches), 'eager1': run(fit_host_batches), 'eager2': run(fit_host_batches),
            'graph0': run(fit_host_batches_graph), 'graph1': run(fit_host_batches_graph)}
    names = list(runs)
    print('losses eager0', ['%.6f' % x for x in runs['eager0'][0]])
    print('losses graph0', ['%.6f' % x for x in runs['graph0'][0]])
    for i, a in enumerate(names):
        for b_ in names[i + 1:]:
            la, pa = runs[a]
            lb, pb = runs[b_]
            dl = max(abs(x - y) for x, y in zip(la, lb))
            print(f'(a/b) {a} vs {b_}: max loss diff {dl:.3e} (rel {dl / max(la):.3e})  param rel {rel(pb, pa):.3e}  '
                  f'n params differing > 0.5 lr: {int(((pa - pb).abs() > 0.5 * args.lr).sum())}')

    # (c) optimizer alone with identical gradients
    a = RenderTrainer(dev, seed=2, warmup_steps=3, target_lr=args.lr)
    b_ = RenderTrainer(dev, seed=2, warmup_steps=3, target_lr=args.lr)
    g = torch.Generator(device='cuda').manual_seed(0)
    worst = 0.0
    for i in range(8):
        grad = torch.randn(a.params.shape, generator=g, device=dev) * (3.0 if i == 2 else 0.1)
        a.params.grad, b_.params.grad = grad.clone(), grad.clone()
        a.apply_gradients()
        b_.apply_gradients_on_device()
        assert float(b_._t_dev) == a.iterations
        worst = max(worst, rel(b_.params, a.params))
    print(f'(c) device-side vs host-side Adam, same gradients, 8 steps on the GPU: worst param rel diff {worst:.3e}')

    # (d) the device-side update replayed from a graph
    a = RenderTrainer(dev, seed=2, warmup_steps=3, target_lr=args.lr)
    b_ = RenderTrainer(dev, seed=2, warmup_steps=3, target_lr=args.lr)
    gbuf = torch.zeros_like(b_.params)
    b_.params.grad = gbuf
    b_._t_dev = torch.zeros((), dtype=torch.float64, device=dev)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        pass
    graph = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(graph):
        b_.apply_gradients_on_device()
    worst = 0.0
    for i in range(8):
        grad = torch.randn(a.params.shape, generator=g, device=dev) * (3.0 if i == 2 else 0.1)
        a.params.grad = grad.clone()
        gbuf.copy_(grad)
        a.apply_gradients()
        graph.replay()
        assert float(b_._t_dev) == a.iterations, (float(b_._t_dev), a.iterations)
        worst = max(worst, rel(b_.params, a.params))
    print(f'(d) graph-replayed device-side Adam vs host-side Adam: worst param rel diff {worst:.3e}')

    # (e) gradient of a graph replay vs an eager backward at the same parameters (lr = 0 keeps them equal)
    tr = RenderTrainer(dev, seed=3, bias_std=0.02, warmup_steps=4, target_lr=0.0)
    dbatch = [t.to(dev) for t in batches[0]]
    f = feats.clone().requires_grad_(True)
    tr.train_step(dbatch[:5], f, dbatch[5], dbatch[6], dbatch[7], NEAR, FAR)
    g_eager = [None, None]
    for k in range(2):
        tr.loss_and_backward(dbatch[:5], f, dbatch[5], dbatch[6], dbatch[7], NEAR, FAR)
        g_eager[k] = tr.params.grad.clone()
    replay = tr.capture(dbatch[:5], f, dbatch[5], dbatch[6], dbatch[7], NEAR, FAR)
    g_graph_t = tr.params.grad                  # graph-owned gradient buffer (clamped in place)
    replay()
    torch.cuda.synchronize()
    print(f'(e) gradient eager vs eager {rel(g_eager[1], g_eager[0]):.3e}; graph vs eager '
          f'{rel(g_graph_t, g_eager[0].clamp(-1, 1)):.3e}; |g| max {float(g_eager[0].abs().max()):.3e}; '
          f'fraction of |g| < 1e-6 max: {float((g_eager[0].abs() < 1e-6 * g_eager[0].abs().max()).float().mean()):.3f}; '
          f'sign flips eager/eager: {int((torch.sign(g_eager[0]) != torch.sign(g_eager[1])).sum())}')


if __name__ == '__main__':
    main()
