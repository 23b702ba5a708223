#!/usr/bin/env python
"""Benchmark of the CLIP-NeRF render hot path (BASELINE.json metric: rays/s, forward + backward).

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port)

Workload (config 2 of BASELINE.json): one scene per GPU, 3 source views of 480x640 with 256-channel
fp32 feature maps, 4096 rays x (64 coarse + 128 fine) samples, loss MSE(coarse)+MSE(fine), backward
including the feature-map gradient, per-element gradient clip and Adam on both NeRF MLPs.  Rays are
sharded over GPUs (weak scaling: every rank renders its own 4096 rays); the only collective is the
all-reduce of the flat NeRF gradient buffer.  Synthetic data, random-init weights.

One JSON line is printed by rank 0 (see the driver contract in the task description)."""
from __future__ import annotations

import argparse
import contextlib
import gc
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'thesis-clip-nerf_b200'))

import torch  # noqa: E402

NEAR, FAR = 0.3, 1.3
H, W, C = 480, 640, 256
FLOP_PER_RAY = {1: 282.97e6, 3: 621.23e6}            # fwd+bwd = 3 x fwd, BASELINE.md section 3
GATHER_BYTES_PER_ROW = 4 * 259 * 4                    # 4 corners x 259 channels x fp32


def _peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            p = json.load(f)
        return dict(hbm_gbs=float(p['hbm_gbs']), tf_burst=float(p['bf16_tflops']),
                    tf_sustained=float(p.get('bf16_tflops_sustained', p['bf16_tflops'])), source='measured')
    except Exception:
        return dict(hbm_gbs=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source='fallback')


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the GPU is under load.  The sampler is started BEFORE the
    warm-up steps (nvidia-smi needs a few hundred ms to produce its first line) and polls every 20 ms; `mark()` brackets
    the timed region, `summary()` uses the samples received inside it and -- when the region was shorter than a polling
    period -- the samples of the last second of load before its end (warm-up steps run the same kernels)."""

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index
        self.t0 = self.t1 = None

    def __enter__(self):
        q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
             'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
             'clocks_event_reasons.sw_power_cap')
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.gpu), f'--query-gpu={q}',
                                          '--format=csv,noheader,nounits', '-lms', '20'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(',')]))

    def mark(self, start: bool):
        if start:
            self.t0 = time.time()
        else:
            self.t1 = time.time()

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.05)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        t0 = self.t0 if self.t0 is not None else 0.0
        t1 = self.t1 if self.t1 is not None else time.time()
        inside = [r for ts, r in self.rows if t0 <= ts <= t1 + 0.03]
        window = 'timed region'
        if len(inside) < 2:
            inside = [r for ts, r in self.rows if t1 - 1.0 <= ts <= t1 + 0.03]
            window = 'last second of load up to the end of the timed region (warm-up + timed steps)'
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in inside:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, val in zip(names, r[3:7]):
                if val.lower().startswith('active'):
                    reasons.add(n)
        if not sm:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['unavailable'])
        sm.sort()
        return dict(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm), window=window)


def make_scene(seed, nr, nv, device=None, pinned=False):
    """synthetic scene of SURVEY.md 8(d): hemisphere cameras, random target-view pixels."""
    from clipnerf_b200 import synthetic
    return synthetic.scene(seed, nr, nv, H, W, C, device=device, pin=pinned)


# ----------------------------------------------------------------------------------------------
# reference arm: the reference's algorithm on the host cores (oracle port; TensorFlow is absent)
# ----------------------------------------------------------------------------------------------
def oracle_step(sc, params, nr_sample):
    from oracle import nerf_oracle as O
    ro, rd, img, K, E = sc['inputs']
    inputs = (ro[:, :nr_sample], rd[:, :nr_sample], img, K, E)
    p = {k: v.detach().requires_grad_(True) for k, v in params.items()}
    feats = sc['features'].detach().requires_grad_(True)
    out = O.mvnerf_call(inputs, feats, p, near=NEAR, far=FAR, n_samples=64, u_strat=sc['u_strat'][:, :nr_sample],
                        u_pdf=sc['u_pdf'][:, :nr_sample])
    loss = O.train_loss(out, sc['target'][:, :nr_sample])
    loss.backward()
    return float(loss)


def cpu_baseline(nv, nr_sample, repeats=1):
    from oracle import nerf_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sc = make_scene(0, max(nr_sample, 2), nv)
    params = O.init_params(0)
    oracle_step(sc, params, min(16, nr_sample))          # warm-up (allocator, thread pool)
    times = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        oracle_step(sc, params, nr_sample)
        times.append(time.perf_counter() - t0)
    t = sorted(times)[len(times) // 2]
    return dict(value=nr_sample / t, unit='rays/s', cores=cores, kind='port',
                sample=f'{nr_sample} rays of the workload (same scene shape, fwd+bwd incl. feature-map gradient), '
                       f'oracle port of the reference TF path, torch CPU fp32, {t:.1f} s per pass'), t


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    nr_sample = args.ref_rays
    base, _ = cpu_baseline(args.nv, nr_sample)
    times = []
    from oracle import nerf_oracle as O
    sc = make_scene(0, nr_sample, args.nv)
    params = O.init_params(0)
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        oracle_step(sc, params, nr_sample)
        if i >= args.warmup:
            times.append(time.perf_counter() - t0)
    ms = 1e3 * sum(times) / len(times)
    value = nr_sample / (ms / 1e3)
    base['value'] = value
    print(json.dumps({
        'impl': 'reference', 'metric': 'rays_per_sec_fwd_bwd', 'value': value, 'unit': 'rays/s', 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': workload_config(args, args.rays),          # this repo's arm's config; each step = the bounded sample below
        'cpu_baseline': base,
        'e2e': {'value': value, 'unit': 'rays/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }))


def workload_config(args, nr):
    tag = 'cfg2' if (nr, args.nv) == (4096, 3) else 'cfg3(i)' if (nr, args.nv) == (8192, 1) else \
        'cfg5 per-GPU shard' if nr >= 131072 else 'custom'
    return {'workload': f'{tag}: CLIP-feature-conditioned NeRF render fwd+bwd, {args.nv} views {H}x{W}x{C} fp32 maps, '
                        f'{nr} rays x (64+128) samples per GPU, MSE(coarse)+MSE(fine), feature-map gradient, clip+Adam',
            'rays_per_gpu': nr, 'n_views': args.nv, 'samples': '64+128', 'image': [H, W], 'feature_channels': C,
            'parallelism': f'ray-sharded dp{args.gpus}', 'l2': 'inputs larger than L2 (943 MB of feature maps per scene)'}


# ----------------------------------------------------------------------------------------------
# this repo's arm
# ----------------------------------------------------------------------------------------------
def run_b200(args):
    import torch.distributed as dist
    from clipnerf_b200 import _abi
    from clipnerf_b200.trainer import RenderTrainer, fit_host_batches, fit_host_batches_graph

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    assert world == args.gpus, f'--gpus {args.gpus} but WORLD_SIZE={world} (launch with torch.distributed.run)'
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        # NCCL's own log lines (version banner, NCCL_DEBUG=INFO) go to stderr: stdout carries the one JSON line
        os.environ.setdefault('NCCL_DEBUG_FILE', '/dev/stderr')
        dist.init_process_group('nccl', device_id=dev)
    _abi.load()

    nr, nv = args.rays, args.nv
    sc = make_scene(100 + rank, nr, nv, device=dev, pinned=True)
    feats = sc['features'].requires_grad_(True)
    trainer = RenderTrainer(dev, seed=0)
    dev_inputs = [t.to(dev, non_blocking=True) for t in sc['inputs']]
    dev_u = [sc[k].to(dev, non_blocking=True) for k in ('u_strat', 'u_pdf', 'target')]

    def step_resident():
        feats.grad = None
        return trainer.train_step(dev_inputs, feats, dev_u[2], dev_u[0], dev_u[1], NEAR, FAR, args.chunk_rays)

    h2d_bytes = sum(t.numel() * t.element_size() for t in sc['inputs']) + \
        sum(sc[k].numel() * sc[k].element_size() for k in ('u_strat', 'u_pdf', 'target'))

    use_graph = nr <= args.chunk_rays and not args.no_graph      # (ray-chunked batches launch eagerly)
    host_batch = list(sc['inputs']) + [sc['target'], sc['u_strat'], sc['u_pdf']]
    assert all(t.is_pinned() for t in host_batch)

    def run_e2e(steps):
        # the call a user makes: a training loop over host batches (the reference's 5-tuple + targets + random
        # draws, pinned memory).  Every step uploads its own batch (copy stream, one batch ahead of the kernels)
        # and returns its loss to the host.
        if use_graph:
            # the training step replayed from CUDA graphs (un-chunked batch): per step the host enqueues the
            # uploads, one graph launch and the loss read-back, so a slow host no longer bounds the loop
            losses = fit_host_batches_graph(trainer, (host_batch for _ in range(steps)), feats, NEAR, FAR)
        else:
            losses = fit_host_batches(trainer, (host_batch for _ in range(steps)), feats, NEAR, FAR, args.chunk_rays)
        assert len(losses) == steps and all(l == l for l in losses)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, loop=False):
        # the cyclic collector is run before and held during the timed region (as timeit does): a generation-2 pass
        # over the autograd garbage of earlier steps can take longer than the two steps the launch queue runs ahead
        gc.collect()
        gc.disable()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        try:
            if loop:
                fn(steps)
            else:
                for _ in range(steps):
                    fn()
        finally:
            gc.enable()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / steps

    with ClockSampler(local) as clocks:
        n_warm = max(args.warmup, 3)
        launches0 = _abi.LAUNCHES
        for _ in range(n_warm):
            step_resident()
        launches_per_step = (_abi.LAUNCHES - launches0) // n_warm
        step_timed = step_resident
        if use_graph:
            # un-chunked batch: the step (this library's kernel launches, loss, NCCL gradient all-reduce on > 1 rank,
            # optimizer) is recorded into a CUDA graph once and replayed -- the same launches without the host's per-launch cost between them
            step_timed = trainer.capture(dev_inputs, feats, dev_u[2], dev_u[0], dev_u[1], NEAR, FAR)
            for _ in range(2):
                step_timed()
        torch.cuda.synchronize()
        clocks.mark(True)
        ms = timed(step_timed, args.steps)
        clocks.mark(False)
    launches = launches_per_step * args.steps      # kernels of this library inside the timed region (replayed or eager)
    run_e2e(3)          # warm the loop's own allocations (device batch slots, pinned loss ring, copy stream)
    ms_e2e = timed(run_e2e, args.steps, loop=True)

    # ---- per-stage device times (CUDA events around every ABI call), separate instrumented pass
    stages = {}
    records = []

    @contextlib.contextmanager
    def timer(name):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        yield
        b.record()
        records.append((name, a, b))

    _abi.TIMER = timer
    n_prof = min(args.steps, 3)
    for _ in range(n_prof):
        step_resident()
    torch.cuda.synchronize()
    _abi.TIMER = None
    for name, a, b in records:
        st = stages.setdefault(name, [0.0, 0])
        st[0] += a.elapsed_time(b)
        st[1] += 1
    stage_ms = {k: v[0] / n_prof for k, v in stages.items()}            # ms per step
    stage_calls = {k: v[1] // n_prof for k, v in stages.items()}

    peaks = _peaks()
    rows_per_step = nr * 192 * nv
    # algorithmic work per step of each stage: gather reads 4144 B per sample-view row in the
    # forward; the backward re-reads them for the fine pass (coordinate gradient) and scatter-adds
    # (read+write) for both passes
    alg = {
        'cn_project_gather_fwd': ('hbm', rows_per_step * GATHER_BYTES_PER_ROW),
        'cn_project_gather_bwd': ('hbm', rows_per_step * GATHER_BYTES_PER_ROW * 2 + nr * 128 * nv * GATHER_BYTES_PER_ROW),
        'cn_mlp_feature_fwd': ('tensor', nr * 192 * nv * 293632.0),
        'cn_mlp_fusion_fwd': ('tensor', nr * 192 * 197632.0),
        'cn_mlp_block_bwd': ('tensor', 2.0 * (nr * 192 * nv * 3 + nr * 192 * 3) * 4 * 128 * 128),
        'cn_mlp_l0_dgrad': ('tensor', nr * 192 * nv * 2.0 * 384 * 128),
        'cn_mlp_l0_wgrad': ('tensor', nr * 192 * nv * 2.0 * 384 * 128),
    }
    # measured DRAM traffic of every hot kernel (ncu capture of this command, cfg2 only): actual bytes moved through
    # HBM per step / live stage time = the DRAM bandwidth the stage really draws (the `achieved_gbs` of the gather
    # stages counts ALGORITHMIC bytes, most of which are served by L2 after the spatial binning -- hence > peak)
    ncu_traffic = {}
    try:
        if (nr, nv) == (4096, 3):
            with open(os.path.join(ROOT, 'profiles', 'r01_v8_traffic.json')) as f:
                raw = json.load(f)
            for k, v in raw.items():
                if isinstance(v, dict):
                    ncu_traffic[k.split(' ')[0]] = v['dram_bytes_per_launch'] * v['launches_per_step']
    except Exception:
        ncu_traffic = {}
    stage_report = {}
    for k, t in sorted(stage_ms.items(), key=lambda kv: -kv[1]):
        ent = {'ms_per_step': round(t, 4), 'calls_per_step': stage_calls[k]}
        if k in ncu_traffic:
            ent['dram_gbs_ncu_bytes'] = round(ncu_traffic[k] / (t * 1e-3) / 1e9, 1)
            ent['dram_frac'] = round(ncu_traffic[k] / (t * 1e-3) / 1e9 / peaks['hbm_gbs'], 4)
        if k in alg:
            bound, work = alg[k]
            if bound == 'hbm':
                ent.update(bound='hbm', achieved_gbs=round(work / (t * 1e-3) / 1e9, 1),
                           frac=round(work / (t * 1e-3) / 1e9 / peaks['hbm_gbs'], 4))
            else:
                ent.update(bound='tensor', achieved_tflops=round(work / (t * 1e-3) / 1e12, 1),
                           frac=round(work / (t * 1e-3) / 1e12 / peaks['tf_sustained'], 4))
        stage_report[k] = ent
    dom = next(iter(stage_report))
    d = stage_report[dom]
    # DRAM traffic per launch of the dominant kernel from the committed ncu --set full capture of this command
    # (cfg2 only: the capture is of the default workload)
    traffic = None
    try:
        if (nr, nv) == (4096, 3):
            with open(os.path.join(ROOT, 'profiles', 'r01_v8_traffic.json')) as f:
                traffic = json.load(f).get(dom, {}).get('dram_bytes_per_launch')
    except Exception:
        traffic = None
    if d.get('bound') == 'hbm':
        roofline = {'kernel': dom, 'bound': 'hbm', 'achieved': d['achieved_gbs'], 'peak': peaks['hbm_gbs'],
                    'unit': 'GB/s', 'frac': d['frac'], 'traffic': traffic, 'peak_source': peaks['source']}
    elif d.get('bound') == 'tensor':
        roofline = {'kernel': dom, 'bound': 'tensor', 'achieved': d['achieved_tflops'], 'peak': peaks['tf_sustained'],
                    'unit': 'TFLOP/s', 'frac': d['frac'], 'traffic': traffic, 'traffic_unit': 'bytes per launch (ncu dram read+write, profiles/r01_v8_traffic.json)',
                    'peak_source': peaks['source'] + ' (sustained)'}
    else:
        roofline = {'kernel': dom, 'bound': 'hbm', 'achieved': None, 'peak': peaks['hbm_gbs'], 'unit': 'GB/s',
                    'frac': None, 'traffic': None, 'peak_source': peaks['source']}
    mlp_ms = sum(t for k, t in stage_ms.items() if k.startswith('cn_mlp'))
    whole = {
        'mlp_tensor_frac_of_sustained': round(nr * FLOP_PER_RAY.get(nv, 0) / (mlp_ms * 1e-3) / 1e12 / peaks['tf_sustained'], 4)
        if mlp_ms > 0 and nv in FLOP_PER_RAY else None,
        'mlp_ms_per_step': round(mlp_ms, 4),
    }

    if rank == 0:
        value = world * nr / (ms * 1e-3)
        out = {
            'metric': 'rays_per_sec_fwd_bwd', 'value': value, 'unit': 'rays/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': max(args.warmup, 3), 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'bf16 operands / f32 accumulate (f32 maps, sampling, compositing)',
            'data': 'synthetic', 'config': workload_config(args, nr),
            'e2e': {'value': world * nr / (ms_e2e * 1e-3), 'unit': 'rays/s', 'ms_per_step': ms_e2e,
                    'h2d_bytes_per_step': h2d_bytes, 'd2h_bytes_per_step': 4,
                    'loop': 'fit_host_batches_graph (train step replayed from a CUDA graph)' if use_graph else 'fit_host_batches (eager launches)',
                    'note': 'host batch = the reference 5-tuple (rays, images, K, E^-1) + targets + '
                            'random draws in pinned memory, uploaded every step on a copy stream one batch ahead; loss '
                            'read back every step; combined_features stay on the device, where the upstream encoders '
                            'produce them'},
            'gpu_launches': launches, 'launch_mode': 'cuda graph replay' if use_graph else 'eager', 'clocks': clocks.summary(), 'roofline': roofline, 'stages': stage_report,
            'mlp': whole,
        }
        if world == 1 and not args.no_cpu_baseline:
            base, _ = cpu_baseline(nv, args.ref_rays)
            out['cpu_baseline'] = base
        print(json.dumps(out))
    if world > 1:
        sys.stdout.flush()
        if use_graph:
            # graphs that recorded NCCL work must be gone before the communicator is torn down
            trainer.__dict__.pop('_graphs', None)
            step_timed = None  # noqa: F841
            gc.collect()
            torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--graph-multi', action='store_true', help='(default now) replay CUDA graphs, with the NCCL all-reduce recorded, on > 1 rank')
    ap.add_argument('--no-graph', action='store_true', help='e2e loop with eager launches instead of CUDA-graph replay')
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--rays', type=int, default=4096, help='rays per GPU per step')
    ap.add_argument('--nv', type=int, default=3, help='source views')
    ap.add_argument('--ref-rays', type=int, default=512, help='bounded CPU sample (rays) of the reference arm')
    ap.add_argument('--chunk-rays', type=int, default=16384,
                    help='batches with more rays per GPU are rendered in chunks of this many rays (gradients accumulate)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
