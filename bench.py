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
# workloads (BASELINE.json configs)
# ----------------------------------------------------------------------------------------------
WORKLOADS = {
    # name: (rays per GPU, source views, what BASELINE.json calls it)
    'cfg2': (4096, 3, 'cfg2: CLIP-feature-conditioned NeRF render fwd+bwd, 3 views, 4096 rays x (64+128) samples'),
    'cfg1': (1024, 1, 'cfg1: train_nerf.py default config (nerf_1_view), 1024 rays'),
    'cfg3i': (8192, 1, 'cfg3(i): NeRF training step, 8192 rays/GPU, 1 view, ray-sharded DDP'),
    'cfg5shard': (131072, 3, 'cfg5 per-GPU shard: 1M rays x 192 samples over 8 GPUs = 131072 rays/GPU, ray-chunked'),
}


def workload_config(args, nr):
    tag = WORKLOADS[args.workload][2] if args.workload in WORKLOADS and WORKLOADS[args.workload][:2] == (nr, args.nv) \
        else f'custom: {nr} rays, {args.nv} views'
    low = getattr(args, 'lowres_features', False)
    maps = (f'{H // 2}x{W // 2}x{C} fp32 maps = the input of the final 2x bilinear UpSampling2D, upsample folded into the '
            f'gather (same function of the same data as the {H}x{W} maps)') if low else f'{H}x{W}x{C} fp32 maps'
    cfg = {'workload': f'{tag}; {args.nv} views {maps}, {nr} rays x (64+128) samples per GPU, '
                       f'MSE(coarse)+MSE(fine), feature-map gradient, clip+Adam',
           'rays_per_gpu': nr, 'n_views': args.nv, 'samples': '64+128', 'image': [H, W], 'feature_channels': C,
           'parallelism': f'ray-sharded dp{args.gpus}',
           'l2': ('L2 flushed by the step itself (every step streams > 10 GB of activations through HBM; the 236 MB of '
                  'low-resolution maps exceed the 126 MB L2 as well)') if low
           else 'inputs larger than L2 (943 MB of feature maps per scene)'}
    if low:
        cfg['lowres_features'] = True
    return cfg


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
    nr_sample = min(args.ref_rays, args.rays)
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
    cfg = workload_config(args, args.rays)       # this repo's arm's workload ...
    # ... of which every CPU pass renders a BOUNDED SAMPLE: rays_per_pass rays (not rays_per_gpu), same maps / views
    cfg['rays_per_pass'] = nr_sample
    cfg['reference_arm'] = (f'one host process on {base["cores"]} cores, whatever --gpus says; each pass = {nr_sample} of the '
                            f'{args.rays} rays of the workload on the full-size maps (the fixed per-pass cost of the 0.94 GB '
                            f'map -- zero fill + gradient -- is spread over {nr_sample} rays)')
    print(json.dumps({
        'impl': 'reference', 'metric': 'rays_per_sec_fwd_bwd', 'value': value, 'unit': 'rays/s', 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic', 'config': cfg, 'cpu_baseline': base,
        'host_processes': 1,
        'e2e': {'value': value, 'unit': 'rays/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }))


# ----------------------------------------------------------------------------------------------
# this repo's arm
# ----------------------------------------------------------------------------------------------
def _traffic_file(lowres=False):
    for name in (('r02_lowres_traffic.json',) if lowres else ('r02_traffic.json', 'r01_v8_traffic.json')):
        path = os.path.join(ROOT, 'profiles', name)
        if os.path.exists(path):
            return path
    return None


def run_b200(args):
    import numpy as np
    import torch.distributed as dist
    from clipnerf_b200 import _abi
    from clipnerf_b200.synthetic import SyntheticNerfDataset
    from src.lib.data_generator.mvnerf import MVNeRFDataGenerator
    from src.lib.mvnerf.model_v0 import MVVNeRFRenderer

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
    # combined_features: produced upstream by the image encoders (out of scope), resident on the device; the plug-in
    # below hands them to the model and receives their gradient every step (model.feature_fn)
    Hf, Wf = (H // 2, W // 2) if args.lowres_features else (H, W)
    if args.lowres_features:
        # the input of the encoder's final 2x bilinear UpSampling2D (layers.py:589,659): the upsample is folded into the
        # gather (CnScene.feat_up = 2), the map and its gradient stay 240 x 320
        sc['features'] = sc['features'][:, :, ::2, ::2].contiguous()
    maps = torch.nn.Parameter(sc['features'].reshape(nv, Hf, Wf, C))
    model = MVVNeRFRenderer(n_rays_train=nr, n_rays_infer=nr, n_views=nv, near=NEAR, far=FAR,
                            original_image_size=(H, W), feature_fn=lambda images: maps, device=dev)
    torch.manual_seed(0)
    model.compile()                      # default optimizer: RenderTrainer on the model's flat parameter buffer
    model.max_rays_per_step = args.chunk_rays
    trainer = model.optimizer
    trainer.overlap_allreduce = not args.no_overlap_allreduce   # all-reduce started inside the backward, under the map sweep
    feats = maps.detach().reshape(1, nv, Hf, Wf, C).requires_grad_(True)       # same storage, the trainer's leaf
    dev_inputs = [t.to(dev, non_blocking=True) for t in sc['inputs']]
    dev_u = [sc[k].to(dev, non_blocking=True) for k in ('u_strat', 'u_pdf', 'target')]

    def step_resident():
        feats.grad = None
        return trainer.train_step(dev_inputs, feats, dev_u[2], dev_u[0], dev_u[1], NEAR, FAR, args.chunk_rays)

    use_graph = nr <= args.chunk_rays and not args.no_graph      # (ray-chunked batches launch eagerly)

    # ---- end to end: MVVNeRFRenderer.fit over a MVNeRFDataGenerator (what src/train_nerf.py:52 calls)
    def make_generator(n_scenes, seed):
        ds = SyntheticNerfDataset(n_scenes=n_scenes, n_perspectives=nv + 1, H=H, W=W, seed=seed)
        return MVNeRFDataGenerator(ds, n_rays_train=nr, batch_size=1, n_views=nv, shuffle=True)
    np.random.seed(1000 + rank)
    gen_warm, gen_timed = make_generator(8, 7 + rank), make_generator(args.steps, 11 + rank)   # 8 warm batches: every slot of the pinned ring (6) exists
    item = gen_warm.compact_item(0)
    h2d_bytes = sum(int(np.asarray(a).nbytes) for a in item[0]) + int(np.asarray(item[1]).size) * 4   # targets go up as fp32
    n_tracked = [0]

    def run_e2e(steps):
        gen = gen_timed if steps == args.steps else gen_warm
        n0 = len(model.loss_tracker)
        hist = model.fit(gen, epochs=1, graph=use_graph)
        n_tracked[0] = len(model.loss_tracker) - n0
        assert n_tracked[0] == len(gen) and all(l == l for l in hist['loss'])

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
        mine = e0.elapsed_time(e1) / steps
        ms = torch.tensor([mine], device=dev)
        per_rank = [mine]
        if world > 1:
            allr = [torch.zeros_like(ms) for _ in range(world)]
            dist.all_gather(allr, ms)
            per_rank = [float(t) for t in allr]
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms), per_rank

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
        ms, ms_per_rank = timed(step_timed, args.steps)
        clocks.mark(False)
    launches = launches_per_step * args.steps      # kernels of this library inside the timed region (replayed or eager)
    # > 1 rank: the same step WITHOUT the gradient all-reduce (every rank on its own), to attribute the growth of the
    # step from 1 to N ranks: collective (+ the rank skew it absorbs) vs per-GPU slow-down on a fully loaded node
    ms_solo_per_rank = None
    if world > 1 and use_graph:
        trainer.world = 1
        solo = trainer.capture(dev_inputs, feats, dev_u[2], dev_u[0], dev_u[1], NEAR, FAR)
        trainer.world = world
        for _ in range(2):
            solo()
        _, ms_solo_per_rank = timed(solo, args.steps)
        del solo
    run_e2e(8)          # warm the loop: device batch slots, all 6 pinned ring slots, copy stream, one graph per buffer set
    ms_e2e, ms_e2e_per_rank = timed(run_e2e, args.steps, loop=True)

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
    real_allreduce = trainer.allreduce_gradients

    def timed_allreduce():
        with timer('nccl_allreduce_gradients'):
            real_allreduce()
    if world > 1:
        trainer.allreduce_gradients = timed_allreduce
    for _ in range(n_prof):
        step_resident()
    torch.cuda.synchronize()
    _abi.TIMER = None
    trainer.allreduce_gradients = real_allreduce
    for name, a, b in records:
        st = stages.setdefault(name, [0.0, 0])
        st[0] += a.elapsed_time(b)
        st[1] += 1
    stage_ms = {k: v[0] / n_prof for k, v in stages.items()}            # ms per step
    stage_calls = {k: v[1] // n_prof for k, v in stages.items()}

    peaks = _peaks()
    rows_per_step = nr * 192 * nv
    # algorithmic work per step of each stage.  Tensor stages: credited FLOPs (recompute not
    # counted).  Gather stages: SURVEY 8(d)'s 4 144 B per sample-view row -- after the spatial binning most of these
    # bytes are served from shared memory, so they are reported as `algorithmic_gbs` only; the roofline fraction of a
    # gather stage is the DRAM traffic ncu measured for it (profiles/*_traffic.json) over its live time.
    alg = {
        'cn_project_gather_fwd': ('hbm', rows_per_step * GATHER_BYTES_PER_ROW),
        'cn_project_gather_bwd': ('hbm', nr * 128 * nv * GATHER_BYTES_PER_ROW),
        'cn_feature_grad': ('hbm', rows_per_step * GATHER_BYTES_PER_ROW * 2),
        'cn_mlp_feature_fwd': ('tensor', nr * 192 * nv * 293632.0),
        'cn_mlp_fusion_fwd': ('tensor', nr * 192 * 197632.0),
        'cn_mlp_block_bwd': ('tensor', 2.0 * (nr * 192 * nv * 3 + nr * 192 * 3) * 4 * 128 * 128),
        'cn_mlp_l0_dgrad': ('tensor', nr * 192 * nv * 2.0 * 384 * 128),
        'cn_mlp_l0_wgrad': ('tensor', nr * 192 * nv * 2.0 * 384 * 128),
    }
    ncu_traffic, ncu_per_launch, traffic_src = {}, {}, None
    try:
        if (nr, nv) == (4096, 3) and _traffic_file(args.lowres_features):
            traffic_src = os.path.relpath(_traffic_file(args.lowres_features), ROOT)
            with open(_traffic_file(args.lowres_features)) as f:
                raw = json.load(f)
            for k, v in raw.items():
                if isinstance(v, dict) and 'dram_bytes_per_launch' in v:
                    name = k.split(' ')[0]
                    ncu_traffic[name] = ncu_traffic.get(name, 0) + v['dram_bytes_per_launch'] * v['launches_per_step']
                    ncu_per_launch.setdefault(name, v['dram_bytes_per_launch'])
    except Exception:
        ncu_traffic = {}
    stage_report = {}
    for k, t in sorted(stage_ms.items(), key=lambda kv: -kv[1]):
        ent = {'ms_per_step': round(t, 4), 'calls_per_step': stage_calls[k]}
        if k in ncu_traffic:
            ent['dram_gbs'] = round(ncu_traffic[k] / (t * 1e-3) / 1e9, 1)
            ent['dram_frac'] = round(ncu_traffic[k] / (t * 1e-3) / 1e9 / peaks['hbm_gbs'], 4)
        if k in alg:
            bound, work = alg[k]
            if bound == 'hbm':
                ent.update(bound='hbm', algorithmic_gbs=round(work / (t * 1e-3) / 1e9, 1))
                if 'dram_frac' in ent:
                    ent['frac'] = ent['dram_frac']            # roofline fraction = measured DRAM bytes / live time / peak
            else:
                tf = work / (t * 1e-3) / 1e12
                ent.update(bound='tensor', achieved_tflops=round(tf, 1), frac_burst=round(tf / peaks['tf_burst'], 4),
                           frac_sustained=round(tf / peaks['tf_sustained'], 4), frac=round(tf / peaks['tf_burst'], 4))
        stage_report[k] = ent
    dom = next(k for k in stage_report if k in alg)
    d = stage_report[dom]
    traffic = ncu_per_launch.get(dom)
    if d.get('bound') == 'hbm':
        roofline = {'kernel': dom, 'bound': 'hbm', 'achieved': d.get('dram_gbs'), 'peak': peaks['hbm_gbs'],
                    'unit': 'GB/s', 'frac': d.get('frac'), 'traffic': traffic, 'algorithmic_gbs': d['algorithmic_gbs'],
                    'peak_source': peaks['source'], 'traffic_source': traffic_src}
    else:
        # the timed region is a burst (tens of ms at full clocks, no power cap): burst peak is the honest denominator
        roofline = {'kernel': dom, 'bound': 'tensor', 'achieved': d['achieved_tflops'], 'peak': peaks['tf_burst'],
                    'unit': 'TFLOP/s', 'frac': d['frac_burst'], 'frac_burst': d['frac_burst'],
                    'frac_sustained': d['frac_sustained'], 'peak_sustained': peaks['tf_sustained'], 'traffic': traffic,
                    'traffic_unit': f'bytes per launch (ncu dram read+write, {traffic_src})',
                    'peak_source': peaks['source'] + ' (burst)'}
    mlp_ms = sum(t for k, t in stage_ms.items() if k.startswith('cn_mlp'))
    whole = {}
    if nv in FLOP_PER_RAY:
        step_tf = nr * FLOP_PER_RAY[nv] / (ms * 1e-3) / 1e12
        whole = {'step_mlp_tflops': round(step_tf, 1), 'step_frac_of_burst': round(step_tf / peaks['tf_burst'], 4),
                 'step_frac_of_sustained': round(step_tf / peaks['tf_sustained'], 4),
                 'mlp_kernels_ms_per_step': round(mlp_ms, 4),
                 'mlp_kernels_frac_of_burst': round(nr * FLOP_PER_RAY[nv] / (mlp_ms * 1e-3) / 1e12 / peaks['tf_burst'], 4)
                 if mlp_ms > 0 else None}

    # ---- cfg5 strong-scaling shard under the 8-GPU run: ONE scene, its maps broadcast once, 2^20 rays / world per rank
    extra = {}
    if world == 8 and not args.no_cfg5 and (nr, nv) == (4096, 3):
        shard = (1 << 20) // world
        with torch.no_grad():
            dist.broadcast(maps, src=0)                          # features of the one scene, NCCL over NVSwitch
            for t in dev_inputs[2:]:
                dist.broadcast(t, src=0)
        sc5 = make_scene(500 + rank, shard, nv)                  # this rank's ray range of the target view
        in5 = [t.to(dev) for t in sc5['inputs'][:2]] + dev_inputs[2:]
        u5 = [sc5[k].to(dev) for k in ('u_strat', 'u_pdf', 'target')]
        del sc5

        def step5():
            feats.grad = None
            return trainer.train_step(in5, feats, u5[2], u5[0], u5[1], NEAR, FAR, args.chunk_rays)
        step5()
        ms5, ms5_ranks = timed(step5, 3)
        extra['cfg5_strong'] = {'workload': 'cfg5: 2^20 rays x 192 samples of ONE scene (maps broadcast once), sharded '
                                            f'{shard} rays per GPU over {world} GPUs, fwd+bwd, ray-chunked ({args.chunk_rays})',
                                'value': (1 << 20) / (ms5 * 1e-3), 'unit': 'rays/s', 'ms_per_step': ms5,
                                'ms_per_step_per_rank': [round(x, 3) for x in ms5_ranks], 'steps': 3, 'scaling': 'strong'}

    if rank == 0:
        value = world * nr / (ms * 1e-3)
        out = {
            'metric': 'rays_per_sec_fwd_bwd', 'value': value, 'unit': 'rays/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': max(args.warmup, 3), 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'bf16 operands / f32 accumulate (f32 maps, sampling, compositing)',
            'data': 'synthetic', 'config': workload_config(args, nr),
            'e2e': {'value': world * nr / (ms_e2e * 1e-3), 'unit': 'rays/s', 'ms_per_step': ms_e2e,
                    'h2d_bytes_per_step': h2d_bytes, 'd2h_bytes_per_step': 4,
                    'loop': 'MVVNeRFRenderer.fit(MVNeRFDataGenerator) -> ' +
                            ('fit_host_batches_graph (one CUDA-graph launch per step)' if use_graph else 'fit_host_batches (eager, ray-chunked)'),
                    'note': 'the call src/train_nerf.py:52 makes: model.fit over a MVNeRFDataGenerator on a synthetic '
                            'in-memory dataset (its batches assembled by a background thread: random source / target '
                            'views, bbox_biased_sample, get_specific_rays, targets), every step uploads its batch from '
                            'pinned memory (rays, uint8 source images converted on the device, K, E^-1, targets) one batch '
                            'ahead on a copy stream and its loss is read back; the stratified / inverse-CDF draws are made '
                            'on the device as the reference does; combined_features stay on the device, where the upstream '
                            'encoders produce them (model.feature_fn plug-in returns the resident maps and receives their gradient)'},
            'gpu_launches': launches, 'launch_mode': 'cuda graph replay' if use_graph else 'eager',
            'overlap_allreduce': world > 1 and not args.no_overlap_allreduce,
            'clocks': clocks.summary(), 'roofline': roofline, 'stages': stage_report, 'mlp': whole,
        }
        if world > 1:
            out['ms_per_step_per_rank'] = [round(x, 4) for x in ms_per_rank]
            out['e2e']['ms_per_step_per_rank'] = [round(x, 4) for x in ms_e2e_per_rank]
            if ms_solo_per_rank is not None:
                out['ms_per_step_without_allreduce_per_rank'] = [round(x, 4) for x in ms_solo_per_rank]
                out['scaling_note'] = ('with the all-reduce recorded in the graph every rank reports the same step (the '
                                       'collective absorbs the skew); ms_per_step_without_allreduce_per_rank is the same '
                                       'graph without it, all ranks running at once: step minus its max = the cost of the '
                                       'collective, its spread = rank skew, its level vs the 1-GPU run = what a fully '
                                       'loaded node (power cap, shared host) costs every GPU')
        out.update(extra)
        if world == 1 and not args.no_cpu_baseline:
            base, _ = cpu_baseline(nv, min(args.ref_rays, nr))
            out['cpu_baseline'] = base
        if world == 1 and args.other_configs and args.workload == 'cfg2' and not args.lowres_features:
            # this process keeps its buffers; drop what the timed loops held so that the children have the device
            del step_timed, trainer, model, feats, maps
            gc.collect()
            torch.cuda.empty_cache()
            out['other_configs'] = other_configs()
        print(json.dumps(out))
    if world > 1:
        sys.stdout.flush()
        # graphs that recorded NCCL work must be gone before the communicator is torn down
        trainer.__dict__.pop('_graphs', None)
        step_timed = None  # noqa: F841
        gc.collect()
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()


def other_configs():
    """The other BASELINE.json configs, each measured by its own short child process on the same GPU after the main
    measurement (device-timed inside the child; see the tools' docstrings): cfg1 / cfg3(i) / the cfg5 per-GPU shard
    through this file, cfg4 through tools/bench_render_view.py, cfg3(ii) through tools/bench_query.py."""
    py = sys.executable
    runs = {
        'cfg1': [py, os.path.join(ROOT, 'bench.py'), '--workload', 'cfg1', '--steps', '10', '--no-cpu-baseline', '--no-other-configs'],
        'cfg3i': [py, os.path.join(ROOT, 'bench.py'), '--workload', 'cfg3i', '--steps', '10', '--no-cpu-baseline', '--no-other-configs'],
        'cfg2_lowres_features': [py, os.path.join(ROOT, 'bench.py'), '--lowres-features', '--steps', '10', '--no-cpu-baseline'],
        'cfg5_shard_1gpu': [py, os.path.join(ROOT, 'bench.py'), '--workload', 'cfg5shard', '--steps', '2', '--warmup', '3',
                            '--no-cpu-baseline', '--no-other-configs'],
        'cfg4_render_view_224': [py, os.path.join(ROOT, 'tools', 'bench_render_view.py'), '--size', '224x224'],
        'cfg4_render_view_480x640': [py, os.path.join(ROOT, 'tools', 'bench_render_view.py'), '--size', '480x640', '--steps', '3'],
        'cfg3ii_pose_query': [py, os.path.join(ROOT, 'tools', 'bench_query.py'), '--steps', '10'],
    }
    keep = ('case', 'metric', 'value', 'unit', 'ms_per_step', 'ms_per_image', 'rows_per_step', 'gpu_launches', 'launch_mode')
    res = {}
    for name, cmd in runs.items():
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
            lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith('{')]
            if r.returncode != 0 or not lines:
                res[name] = {'error': (r.stderr or 'no output').strip().splitlines()[-1][:200]}
                continue
            ents = []
            for d in lines:
                e = {k: d[k] for k in keep if k in d}
                if isinstance(d.get('config'), dict):
                    e['workload'] = d['config'].get('workload')
                if isinstance(d.get('e2e'), dict):
                    e['e2e'] = {k: d['e2e'][k] for k in ('value', 'unit', 'ms_per_step') if k in d['e2e']}
                ents.append(e)
            res[name] = ents[0] if len(ents) == 1 else ents
        except Exception as ex:          # a child that fails or hangs must not take the headline line with it
            res[name] = {'error': f'{type(ex).__name__}: {ex}'[:200]}
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--no-graph', action='store_true', help='eager launches instead of CUDA-graph replay')
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default='cfg2', choices=sorted(WORKLOADS),
                    help='BASELINE.json config (rays per GPU, views); --rays / --nv override')
    ap.add_argument('--rays', type=int, default=None, help='rays per GPU per step')
    ap.add_argument('--nv', type=int, default=None, help='source views')
    ap.add_argument('--ref-rays', type=int, default=1024, help='bounded CPU sample (rays per pass) of the reference arm')
    ap.add_argument('--chunk-rays', type=int, default=16384,
                    help='batches with more rays per GPU are rendered in chunks of this many rays (gradients accumulate)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--lowres-features', action='store_true',
                    help='feed the 240x320 input of the final 2x bilinear UpSampling2D; the upsample is folded into the gather')
    ap.add_argument('--no-other-configs', dest='other_configs', action='store_false',
                    help='skip the short child runs of the other BASELINE.json configs (cfg1, cfg3, cfg4, cfg5 shard) behind the default line')
    ap.add_argument('--no-overlap-allreduce', action='store_true',
                    help='> 1 rank: all-reduce behind the backward instead of inside it (NCCL stream, under the feature-map sweep)')
    ap.add_argument('--no-cfg5', action='store_true', help='skip the cfg5 strong-scaling shard of the 8-GPU run')
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    args.rays = wl[0] if args.rays is None else args.rays
    args.nv = wl[1] if args.nv is None else args.nv
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
