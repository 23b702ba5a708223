"""where the end-to-end loop loses time against the resident step (diagnostic)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]
import torch
from clipnerf_b200 import synthetic
from clipnerf_b200.trainer import RenderTrainer, fit_host_batches, DevicePrefetcher

dev = torch.device('cuda:0')
sc = synthetic.scene(100, 4096, 3, 480, 640, 256, device=dev, pin=True)
feats = sc['features'].requires_grad_(True)
tr = RenderTrainer(dev, seed=0)
dev_inputs = [t.to(dev) for t in sc['inputs']]
u = [sc[k].to(dev) for k in ('u_strat', 'u_pdf', 'target')]
host_batch = list(sc['inputs']) + [sc['target'], sc['u_strat'], sc['u_pdf']]


def timed(fn, steps=20):
    fn(3)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(steps); e1.record()
    t_host = time.perf_counter() - t0
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, 1e3 * t_host / steps


def resident(n):
    for _ in range(n):
        feats.grad = None
        tr.train_step(dev_inputs, feats, u[2], u[0], u[1], 0.3, 1.3)


def resident_readback(n):
    prev = None
    for _ in range(n):
        feats.grad = None
        loss = tr.train_step(dev_inputs, feats, u[2], u[0], u[1], 0.3, 1.3)
        h = torch.empty((), dtype=torch.float32, pin_memory=True); h.copy_(loss, non_blocking=True)
        ev = torch.cuda.Event(); ev.record()
        if prev is not None:
            prev[1].synchronize(); float(prev[0])
        prev = (h, ev)


def upload_only(n):
    for batch, done in DevicePrefetcher((host_batch for _ in range(n)), dev):
        feats.grad = None
        tr.train_step(batch[:5], feats, batch[5], batch[6], batch[7], 0.3, 1.3)
        done.record(torch.cuda.current_stream(dev))


def e2e(n):
    fit_host_batches(tr, (host_batch for _ in range(n)), feats, 0.3, 1.3)


def upload_alone(n):
    for batch, done in DevicePrefetcher((host_batch for _ in range(n)), dev):
        done.record(torch.cuda.current_stream(dev))


def upload_unused(n):
    """uploads run beside the steps, the steps read the resident inputs"""
    for batch, done in DevicePrefetcher((host_batch for _ in range(n)), dev):
        feats.grad = None
        tr.train_step(dev_inputs, feats, u[2], u[0], u[1], 0.3, 1.3)
        done.record(torch.cuda.current_stream(dev))


def upload_depth3(n):
    for batch, done in DevicePrefetcher((host_batch for _ in range(n)), dev, depth=3):
        feats.grad = None
        tr.train_step(batch[:5], feats, batch[5], batch[6], batch[7], 0.3, 1.3)
        done.record(torch.cuda.current_stream(dev))


def upload_small_only(n):
    """only rays / targets / random draws are uploaded every step; the source images and cameras stay resident"""
    small = [host_batch[0], host_batch[1]] + host_batch[5:]
    for batch, done in DevicePrefetcher((small for _ in range(n)), dev):
        feats.grad = None
        tr.train_step([batch[0], batch[1]] + dev_inputs[2:], feats, batch[2], batch[3], batch[4], 0.3, 1.3)
        done.record(torch.cuda.current_stream(dev))


for name, fn in (('uploads alone', upload_alone), ('uploads beside resident steps', upload_unused),
                 ('prefetched uploads, depth 3', upload_depth3), ('rays-only uploads', upload_small_only)):
    ms, host = timed(fn)
    print(f'{name:40s} device {ms:7.3f} ms/step   host enqueue {host:7.3f} ms/step')

for name, fn in (('resident', resident), ('resident + lagged loss readback', resident_readback),
                 ('prefetched uploads, no readback', upload_only), ('fit_host_batches (e2e)', e2e)):
    ms, host = timed(fn)
    print(f'{name:40s} device {ms:7.3f} ms/step   host enqueue {host:7.3f} ms/step')


def e2e_instrumented(n):
    """fit_host_batches with host timestamps: where does the host spend the step?"""
    cur = torch.cuda.current_stream(dev)
    ring = torch.empty(4, dtype=torch.float32, pin_memory=True)
    pending, k = [], 0
    t_next = t_step = t_sync = 0.0
    it = DevicePrefetcher((host_batch for _ in range(n)), dev)
    while True:
        a = time.perf_counter()
        try:
            batch, done = next(it)
        except StopIteration:
            break
        b = time.perf_counter()
        feats.grad = None
        loss = tr.train_step(batch[:5], feats, batch[5], batch[6], batch[7], 0.3, 1.3)
        done.record(cur)
        h = ring[k % 4]; k += 1
        h.copy_(loss, non_blocking=True)
        ev = torch.cuda.Event(); ev.record(cur)
        pending.append((h, ev))
        c = time.perf_counter()
        if len(pending) > 1:
            hh, e = pending.pop(0)
            e.synchronize(); float(hh)
        d = time.perf_counter()
        t_next += b - a; t_step += c - b; t_sync += d - c
    torch.cuda.synchronize()
    print(f'    host per step: prefetcher {1e3 * t_next / n:.3f} ms, train_step enqueue {1e3 * t_step / n:.3f} ms, '
          f'wait for the previous loss {1e3 * t_sync / n:.3f} ms')


for _ in range(2):
    ms, host = timed(e2e_instrumented, 40)
    print(f'{"e2e instrumented":40s} device {ms:7.3f} ms/step   host {host:7.3f} ms/step')
