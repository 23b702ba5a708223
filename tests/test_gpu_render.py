"""End-to-end parity of the CUDA render path with the oracle's ``MVVNeRFRenderer._call``:
forward outputs (rel-err <= 1e-2) and gradients of MSE(coarse)+MSE(fine) w.r.t. every NeRF
parameter and the feature maps (rel-err <= 2e-2)."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu

NEAR, FAR = 0.3, 1.3


def _run(cuda, sc, named, need_feat_grad=True):
    from clipnerf_b200 import params, render
    pc = params.from_named(named, 'coarse').to(cuda).requires_grad_(True)
    pf = params.from_named(named, 'fine').to(cuda).requires_grad_(True)
    inputs = [t.to(cuda) for t in sc['inputs']]
    feats = sc['features'].to(cuda).requires_grad_(need_feat_grad)
    out = render.render_rays(pc, pf, inputs, feats, sc['u_strat'].to(cuda), sc['u_pdf'].to(cuda), NEAR, FAR)
    return out, pc, pf, feats


def _report(tag, errs, tol_out=1e-2, tol_grad=2e-2, skip=()):
    for k, v in errs.items():
        print(f'[{tag}] {k:62s} {v:.3e}')
    bad = {k: v for k, v in errs.items()
           if k not in skip and not v < (tol_grad if k.startswith('grad') else tol_out)}
    assert not bad, bad


@pytest.mark.parametrize('nv,b', [(1, 1), (3, 1), (2, 2)])
def test_coarse_pass_gradients(cuda, nv, b):
    """loss = MSE(coarse rgb): stratified z is identical on both sides, so every coarse-net
    gradient and the feature-map gradient must agree with the (bf16-operand) oracle."""
    from clipnerf_b200 import params
    from oracle import nerf_oracle as O
    sc = O.synthetic_scene(11, b, nv, 32, 48, 64)
    named = O.init_params(5, bias_std=0.02)
    p = {k: v.clone().requires_grad_(True) for k, v in named.items()}
    feats = sc['features'].clone().requires_grad_(True)
    ref = O.mvnerf_call(sc['inputs'], feats, p, near=NEAR, far=FAR, n_samples=64, u_strat=sc['u_strat'],
                        u_pdf=sc['u_pdf'], quant='bf16')
    (O.mse(sc['target'], ref[0]) + 0.1 * ref[1].mean()).backward()
    out, pc, pf, fd = _run(cuda, sc, named)
    tgt = sc['target'].to(cuda)
    (torch.mean((out[0] - tgt) ** 2) + 0.1 * out[1].mean()).backward()
    errs = {'rgb': rel_err(out[0], ref[0]), 'depth': rel_err(out[1], ref[1])}
    for name, gv in params.to_named(pc.grad.cpu(), 'coarse').items():
        errs['grad ' + name] = rel_err(gv, p[name].grad)
    errs['grad features'] = rel_err(fd.grad, feats.grad)
    assert float(pf.grad.abs().max()) == 0.0
    _report(f'coarse nv={nv}', errs)


@pytest.mark.parametrize('nv', [1, 3])
def test_fine_pass_and_sampler_gradients(cuda, nv):
    """Fine pass + hierarchical sampler with the SAME coarse weights on both sides (so z_fine is
    bit-identical): fine-net gradients, feature-map gradient and the gradient that reaches the
    coarse weights through sort -> inverse CDF (no stop_gradient, model_v0.py:153-158).

    This path is tested with shared weights because, end to end, it is ill-conditioned in the
    reference itself: a 1e-4 relative change of the coarse net moves the fine samples by ~1e-4 m,
    i.e. ~0.1 rad of phase at the top posenc frequency, and changes this gradient by ~20 %."""
    from clipnerf_b200 import ops, params, render
    from oracle import nerf_oracle as O
    sc = O.synthetic_scene(17, 1, nv, 32, 48, 64)
    named = O.init_params(9, bias_std=0.02)
    i_ro, i_rd, img, K, E = sc['inputs']
    g = torch.Generator().manual_seed(3)
    _, z_c = O.sample_along_ray(i_ro, i_rd, NEAR, FAR, 64, sc['u_strat'])
    w_c = (torch.rand(1, 32, 64, generator=g) ** 4 * 0.2)
    # ---- oracle
    p = {k: v.clone().requires_grad_(True) for k, v in named.items()}
    feats = sc['features'].clone().requires_grad_(True)
    w_leaf = w_c.clone().requires_grad_(True)
    zs = O.sample_pdf(0.5 * (z_c[..., 1:] + z_c[..., :-1]), w_leaf[..., 1:-1], 64, sc['u_pdf'])
    all_zs, _ = torch.sort(torch.cat([z_c, zs], -1), dim=-1, stable=True)
    rgb_r, depth_r, _, _, _ = O._render_pass(i_ro, i_rd, all_zs, img * 2 - 1.0, feats, K, E, p, 'fine', nv, 'bf16')
    (O.mse(sc['target'], rgb_r) + 0.1 * depth_r.mean()).backward()
    # ---- CUDA, from the same building blocks render_rays uses
    ro, rd = i_ro.to(cuda), i_rd.to(cuda)
    scene = ops.Scene(img.to(cuda), sc['features'].to(cuda), K.to(cuda), E.to(cuda))
    pf = params.from_named(named, 'fine').to(cuda)
    z_f, perm, _ = ops.sample_pdf_fwd(z_c.to(cuda), w_c.to(cuda), sc['u_pdf'].to(cuda))
    assert torch.equal(z_f.cpu(), all_zs.detach())
    Fp = render._pass_fwd(ops.pack_weights(pf), scene, ro, rd, z_f, True)
    tgt = sc['target'].to(cuda)
    d_rgb = 2.0 * (Fp.rgb - tgt) / tgt.numel()
    d_depth = torch.full_like(Fp.depth, 0.1 / Fp.depth.numel())
    g_f = torch.zeros_like(pf)
    dz_f, dx_f = render._pass_bwd(Fp, pf, g_f, scene, ro, rd, d_rgb, d_depth, None, True, True)
    d_feat = torch.full_like(scene.features, float('nan'))      # every texel is written exactly once
    ops.feature_grad(scene, [(Fp.ws, dx_f, z_f.shape[1], z_f.shape[2])], d_feat, accumulate=False)
    dw_c = ops.sample_pdf_bwd(z_c.to(cuda), w_c.to(cuda), sc['u_pdf'].to(cuda), perm, dz_f)
    errs = {'fine_rgb': rel_err(Fp.rgb, rgb_r), 'fine_depth': rel_err(Fp.depth, depth_r)}
    for name, gv in params.to_named(g_f.cpu(), 'fine').items():
        errs['grad ' + name] = rel_err(gv, p[name].grad)
    errs['grad features'] = rel_err(d_feat, feats.grad)
    errs['grad coarse weights (through the sampler)'] = rel_err(dw_c, w_leaf.grad)
    _report(f'fine nv={nv}', errs)


def test_render_end_to_end(cuda):
    """Whole _call, both losses, against the bf16-operand oracle and the PLAIN-fp32 oracle.

    Outputs (1e-2) and fine-net gradients (2e-2, max-norm) are asserted against both.  The coarse-net and feature-map
    gradients contain the path through the un-stopped sampler, which is ill-conditioned in the reference itself: its
    own TF32 arithmetic moves them by 2 ... 15 % (tests/test_oracle.py::test_reference_gradient_spread_fp32_vs_tf32).
    They are asserted per tensor, statistically, against the fp32 oracle: relative L2 <= 6 x the oracle's own
    fp32-vs-TF32 spread (of that tensor, or the median over the coarse tensors where a tensor's own spread is smaller;
    bf16 operands round 2^3 coarser than TF32; measured 3.5 ... 5 x) and cosine >= 0.8; and in max-norm (6e-2; the map 0.2: single texels do not average over rays) against the oracle
    that rounds operands where the kernels do."""
    from clipnerf_b200 import params
    from oracle import nerf_oracle as O
    from test_oracle import _e2e_gradients, spread
    floor = spread(_e2e_gradients('tf32')[0], _e2e_gradients(None)[0])
    for quant in ('bf16', None):
        sc = O.synthetic_scene(13, 1, 3, 64, 48, 64)
        named = O.init_params(7, bias_std=0.02)
        p = {k: v.clone().requires_grad_(True) for k, v in named.items()}
        feats = sc['features'].clone().requires_grad_(True)
        ref = O.mvnerf_call(sc['inputs'], feats, p, near=NEAR, far=FAR, n_samples=64, u_strat=sc['u_strat'],
                            u_pdf=sc['u_pdf'], quant=quant)
        O.train_loss(ref, sc['target']).backward()
        out, pc, pf, fd = _run(cuda, sc, named)
        tgt = sc['target'].to(cuda)
        (torch.mean((out[0] - tgt) ** 2) + torch.mean((out[2] - tgt) ** 2)).backward()
        errs = {n_: rel_err(a, b_) for n_, a, b_ in zip(('rgb', 'depth', 'fine_rgb', 'fine_depth'), out, ref)}
        got, want = {}, {}
        for net, flat in (('coarse', pc), ('fine', pf)):
            for name, gv in params.to_named(flat.grad.cpu(), net).items():
                errs[f'grad {name}'] = rel_err(gv, p[name].grad)
                got[name], want[name] = gv, p[name].grad
        errs['grad features'] = rel_err(fd.grad, feats.grad)
        got['features'], want['features'] = fd.grad.cpu(), feats.grad
        ill = [k for k in errs if 'coarse' in k or 'features' in k]
        _report(f'e2e quant={quant}', errs, skip=ill)
        if quant == 'bf16':
            bad = {k: errs[k] for k in ill if not errs[k] < (0.2 if 'features' in k else 6e-2)}
            assert not bad, bad
        else:
            sp = spread(got, want)
            med = sorted(v[0] for k, v in floor.items() if k.startswith('coarse'))[len(floor) // 4]
            for k in sorted(sp):
                if k.startswith('coarse') or k == 'features':
                    l2, cos = sp[k]
                    print(f'[e2e vs fp32] {k:58s} rel L2 {l2:.4f} (fp32-vs-TF32 floor {floor[k][0]:.4f}, x{l2 / floor[k][0]:.1f})  '
                          f'cos {cos:.4f} (floor {floor[k][1]:.4f})')
                    assert l2 <= 6.0 * max(floor[k][0], med), (k, l2, floor[k][0], med)
                    assert cos >= 0.8, (k, cos)


def test_render_inference_matches_training_forward(cuda):
    from oracle import nerf_oracle as O
    sc = O.synthetic_scene(12, 2, 2, 16, 32, 40)
    named = O.init_params(6, bias_std=0.02)
    with torch.no_grad():
        out_inf, *_ = _run(cuda, sc, named, need_feat_grad=False)
    out_trn, *_ = _run(cuda, sc, named)
    for a, b_ in zip(out_inf, out_trn):
        assert torch.equal(a, b_.detach())


def test_chained_stage_backward_gives_the_per_block_gradients(cuda, monkeypatch):
    """render.CHAINED_STAGES: the two-launch backward of a pass (cn_mlp_stage_bwd) and the launch-per-block one walk
    the same bf16 g stream; the outputs are bit-identical and the gradients differ by the fp32 atomic order only"""
    from clipnerf_b200 import render
    from oracle import nerf_oracle as O
    sc = O.synthetic_scene(13, 1, 3, 160, 48, 64)
    named = O.init_params(7, bias_std=0.02)
    tgt = sc['target'].to(cuda)
    res = []
    for chained in (False, True):
        monkeypatch.setattr(render, 'CHAINED_STAGES', chained)
        out, pc, pf, fd = _run(cuda, sc, named)
        (torch.mean((out[0] - tgt) ** 2) + torch.mean((out[2] - tgt) ** 2)).backward()
        res.append((out, pc.grad.clone(), pf.grad.clone(), fd.grad.clone()))
    (o0, c0, f0, m0), (o1, c1, f1, m1) = res
    for a, b_ in zip(o0, o1):
        assert torch.equal(a, b_)
    errs = (rel_err(c1, c0), rel_err(f1, f0), rel_err(m1, m0))
    print('chained vs per-block gradients', errs)
    assert errs[0] < 1e-3 and errs[1] < 1e-3 and errs[2] < 3e-3, errs


def test_ray_chunked_training_step_matches_unchunked(cuda):
    """RenderTrainer.loss_and_backward(max_rays=...): gradients accumulated over ray chunks equal the one-shot
    gradients (rays are independent; the loss is normalised by the full batch)."""
    from clipnerf_b200.trainer import RenderTrainer
    from oracle import nerf_oracle as O
    sc = O.synthetic_scene(21, 1, 2, 96, 48, 64)
    inputs = [t.to(cuda) for t in sc['inputs']]
    u_s, u_p, tgt = sc['u_strat'].to(cuda), sc['u_pdf'].to(cuda), sc['target'].to(cuda)
    res = []
    for max_rays in (None, 32):
        tr = RenderTrainer(cuda, seed=3, bias_std=0.02)
        feats = sc['features'].to(cuda).requires_grad_(True)
        loss, out = tr.loss_and_backward(inputs, feats, tgt, u_s, u_p, NEAR, FAR, max_rays)
        res.append((float(loss), tr.params.grad.clone(), feats.grad.clone(), out[2].clone()))
    (l0, g0, f0, o0), (l1, g1, f1, o1) = res
    assert abs(l0 - l1) < 1e-6 * max(1.0, abs(l0))
    assert torch.equal(o0, o1)                                       # per-ray outputs do not depend on the chunking
    # (map gradient: single bf16 rounding flips of the g stream between two runs, see tests/test_gpu_edges.py)
    assert rel_err(g1, g0) < 1e-3 and rel_err(f1, f0) < 3e-3, (rel_err(g1, g0), rel_err(f1, f0))


def _loop_batches(cuda, nr=32):
    from oracle import nerf_oracle as O
    batches = []
    for seed in (41, 42, 43):
        sc = O.synthetic_scene(seed, 1, 2, nr, 48, 64)
        batches.append([t.pin_memory() for t in list(sc['inputs']) + [sc['target'], sc['u_strat'], sc['u_pdf']]])
    feats = O.synthetic_scene(41, 1, 2, nr, 48, 64)['features'].to(cuda)
    return batches, feats


def test_update_kernel_matches_host_adam_on_the_gpu(cuda):
    """cn_adam_step (clip + warm-up schedule + Keras Adam with the step counter on the device) against
    RenderTrainer.apply_gradients (torch ops, host scalars): the SAME gradient tensor per step on both sides, across
    the warm-up boundary, launched eagerly and replayed from a CUDA graph, with and without the readouts."""
    from clipnerf_b200._abi import EMB_PARAMS
    from clipnerf_b200.trainer import RenderTrainer
    for train_readouts in (False, True):
        kw = dict(seed=2, warmup_steps=3, target_lr=1e-3, train_readouts=train_readouts)
        host, eager, graphed = RenderTrainer(cuda, **kw), RenderTrainer(cuda, **kw), RenderTrainer(cuda, **kw)
        gbuf = torch.zeros_like(graphed.params)
        graphed.params.grad = gbuf
        graphed._counter()
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            graphed.apply_gradients_on_device()
        g = torch.Generator(device='cuda').manual_seed(0)
        for i in range(8):
            grad = torch.randn(host.params.shape, generator=g, device=cuda) * (3.0 if i == 2 else 0.1)   # beyond the clip
            host.params.grad, eager.params.grad = grad.clone(), grad.clone()
            gbuf.copy_(grad)
            host.apply_gradients()
            eager._counter()
            eager.apply_gradients_on_device()
            eager.iterations += 1
            graph.replay()
            graphed.iterations += 1
            assert float(eager._t_dev) == float(graphed._t_dev) == host.iterations == i + 1
            for other in (eager, graphed):
                assert rel_err(other.params, host.params) < 1e-6, (i, rel_err(other.params, host.params))
                assert rel_err(other.m, host.m) < 1e-6 and rel_err(other.v, host.v) < 1e-6
            n = eager.n_update                                                 # the updated gradient entries are left clipped
            assert torch.equal(eager.params.grad[:, :n], grad.clamp(-1, 1)[:, :n])
        init = RenderTrainer(cuda, **kw).params.detach()
        same_readouts = torch.equal(graphed.params.detach()[:, EMB_PARAMS:], init[:, EMB_PARAMS:])
        assert same_readouts == (not train_readouts)


def test_graph_replayed_step_walks_the_eager_step(cuda):
    """Per-step lockstep of the CUDA-graph training loop against the eager step, ON THE GPU: before every step of a
    graph-driven trainer its state (parameters, Adam moments, step count) is cloned into a second trainer that takes
    the eager step on the same batch; the two updates must agree, at every step, across the warm-up boundary, with two
    graphs (one per upload buffer set) alternating and an eager step interleaved between replays.  No chaos can build
    up: every comparison starts from identical state."""
    from clipnerf_b200.trainer import RenderTrainer
    batches, feats0 = _loop_batches(cuda)
    kw = dict(seed=3, bias_std=0.02, warmup_steps=4, target_lr=1e-4)
    G = RenderTrainer(cuda, **kw)
    slots = [[t.to(cuda) for t in batches[0]], [t.to(cuda) for t in batches[1]]]
    fg = [feats0.clone().requires_grad_(True) for _ in range(2)]
    graphs = {}
    worst = 0.0
    for i in range(9):
        k = i % 2
        for d, h in zip(slots[k], batches[i % 3]):
            d.copy_(h)
        bt = slots[k]
        E = RenderTrainer(cuda, params=G.params.detach().clone().requires_grad_(True), **{kk: v for kk, v in kw.items()
                                                                                          if kk not in ('seed', 'bias_std')})
        E.m.copy_(G.m), E.v.copy_(G.v)
        E.iterations = G.iterations
        before = G.params.detach().clone()
        if i == 0 or i == 5:                      # first step of a trainer's life, and an eager step between replays
            fg[k].grad = None                     # (autograd accumulates into an existing .grad)
            lg = G.train_step(bt[:5], fg[k], bt[5], bt[6], bt[7], NEAR, FAR)
        else:
            if k not in graphs:
                torch.cuda.synchronize()
                graphs[k] = G.capture(bt[:5], fg[k], bt[5], bt[6], bt[7], NEAR, FAR,
                                      pool=next(iter(graphs.values())).graph.pool() if graphs else None)
            lg = graphs[k]()
        fe = feats0.clone().requires_grad_(True)
        le = E.train_step(bt[:5], fe, bt[5], bt[6], bt[7], NEAR, FAR)
        assert float(G._t_dev) == G.iterations == E.iterations == i + 1          # ONE counter, never stale
        assert abs(float(lg) - float(le)) < 1e-5 * abs(float(le)), (i, float(lg), float(le))
        du_g, du_e = (G.params.detach() - before).double(), (E.params.detach() - before).double()
        # Adam's update is ~lr * sign(g) for the first steps: parameters whose gradient is ~0 may flip on the last
        # bits of the fp32 atomics, everything else must agree -> relative L2 of the whole update
        err = float((du_g - du_e).norm() / du_e.norm())
        worst = max(worst, err)
        assert err < 5e-2, (i, err)
        assert rel_err(fg[k].grad, fe.grad) < 3e-3
    print('graph step vs eager step from identical state: worst relative L2 of the update', worst)


def test_graph_replayed_training_loop_matches_eager(cuda):
    """fit_host_batches_graph (train step replayed from CUDA graphs, step counter / warm-up schedule / Adam bias
    corrections on the device) against the eager loop.  With frozen parameters every replay must return the loss and
    the map gradient of ITS batch.  With training the two loops cannot be bit-equal: the backward's fp32 atomics make
    gradients differ in the last bits from run to run, Adam's first updates are ~lr * sign(g), and the reference's
    un-stopped sampler path (model_v0.py:153-158) amplifies parameter noise.  So the NOISE FLOOR is measured here --
    the eager loop against itself, five runs with identical seeds -- and the graph loop must lie within 2x of it
    (measured on a B200, tools/diag_graph_loop.py: eager/eager loss spread 2e-3 ... 8e-3 relative after 7 steps of 32
    rays, graph/eager 6e-4 ... 9e-3).  The arithmetic of the step itself is pinned without noise by the two tests above."""
    from clipnerf_b200.trainer import RenderTrainer, fit_host_batches, fit_host_batches_graph
    batches, feats = _loop_batches(cuda)
    seq = [batches[i % 3] for i in range(7)]

    def run(fit, lr):
        tr = RenderTrainer(cuda, seed=3, bias_std=0.02, warmup_steps=4, target_lr=lr)
        f = feats.clone().requires_grad_(True)
        losses = fit(tr, iter(seq), f, NEAR, FAR)
        assert tr.iterations == 7 and float(tr._t_dev) == 7.0
        return losses, tr.params.detach().clone(), f.grad.clone()

    # frozen parameters (lr = 0): every replay must return the losses and the map gradient of ITS batch
    (l0, p0, g0), (l1, p1, g1) = run(fit_host_batches, 0.0), run(fit_host_batches_graph, 0.0)
    assert torch.equal(p0, p1)
    print('lr=0: loss diff', max(abs(a - b_) for a, b_ in zip(l0, l1)), 'map grad diff', rel_err(g1, g0))
    assert max(abs(a - b_) for a, b_ in zip(l0, l1)) < 1e-5 * max(l0)
    assert rel_err(g1, g0) < 3e-3, rel_err(g1, g0)
    # training: graph loop within the measured eager-vs-eager floor
    dist = lambda a, b_: max(abs(x - y) for x, y in zip(a[0], b_[0]))  # noqa: E731
    eager = [run(fit_host_batches, 1e-4) for _ in range(5)]
    floor = max(dist(eager[i], eager[j]) for i in range(5) for j in range(i + 1, 5))
    graph = run(fit_host_batches_graph, 1e-4)
    d_graph = min(dist(graph, e) for e in eager)
    print(f'lr=1e-4: eager-vs-eager loss floor {floor:.3e} (5 runs, 10 pairs), graph loop to the nearest eager run {d_graph:.3e}, '
          f'param diff eager/eager {rel_err(eager[1][1], eager[0][1]):.3e} graph/eager {rel_err(graph[1], eager[0][1]):.3e}')
    # The floor of ONE run is itself a sample of ten pairs: on a B200 it came out between 4.9e-4 and 1.7e-3 (absolute, loss
    # ~0.45) over nine runs of this test, the graph loop's distance between 6e-5 and 1.5e-3 -- once 3x a floor that
    # happened to be small.  So the bound is the larger of 2x the measured floor and the largest eager-vs-eager spread
    # seen on this hardware (4e-3 relative); a wrong step counter, bias correction or stale batch is 1e-1 and up.
    assert d_graph <= max(2.0 * floor, 4e-3 * max(eager[0][0])) + 1e-6 * max(eager[0][0]), (d_graph, floor)
    assert abs(graph[0][0] - eager[0][0][0]) < 1e-5 * eager[0][0][0]          # the first step has no history: equal
