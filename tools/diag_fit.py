"""per-step device timeline of MVVNeRFRenderer.fit (end-to-end loop): where does it lose time against the replayed step?"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]
import numpy as np, torch
from clipnerf_b200 import synthetic
from clipnerf_b200.synthetic import SyntheticNerfDataset
from src.lib.data_generator.mvnerf import MVNeRFDataGenerator
from src.lib.mvnerf.model_v0 import MVVNeRFRenderer
H, W, C, nr, nv = 480, 640, 256, 4096, 3
dev = torch.device('cuda:0')
sc = synthetic.scene(100, nr, nv, H, W, C, device=dev, pin=True)
maps = torch.nn.Parameter(sc['features'].reshape(nv, H, W, C))
model = MVVNeRFRenderer(n_rays_train=nr, n_rays_infer=nr, n_views=nv, near=0.3, far=1.3, original_image_size=(H, W),
                        feature_fn=lambda images: maps, device=dev)
model.compile()
def gen(n, seed):
    return MVNeRFDataGenerator(SyntheticNerfDataset(n_scenes=n, n_perspectives=nv + 1, H=H, W=W, seed=seed), n_rays_train=nr,
                               batch_size=1, n_views=nv, shuffle=True)
model.fit(gen(8, 1), epochs=1)
torch.cuda.synchronize()
events, hosts = [], []
orig = model._encoder_backward
def hook(k, feats, f_out):
    orig(k, feats, f_out)
    e = torch.cuda.Event(enable_timing=True); e.record(); events.append(e); hosts.append(time.perf_counter())
model._encoder_backward = hook
g = gen(20, 2)
prod = []
_ci = g.compact_item
def ci(i):
    a = time.perf_counter(); r = _ci(i); prod.append((a, time.perf_counter())); return r
g.compact_item = ci
from clipnerf_b200 import trainer as T
marks = []
def wrap(obj, name, tag):
    f = getattr(obj, name)
    def g_(*a, **k):
        a0 = time.perf_counter(); r = f(*a, **k); marks.append((tag, a0, time.perf_counter())); return r
    setattr(obj, name, g_)
wrap(T.DevicePrefetcher, '_upload', 'upload')
wrap(T, '_step_operands', 'operands')
wrap(T.PinnedBatchQueue, '_produce', 'produce')
wrap(T._LossReader, 'push', 'push')
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
model.fit(g, epochs=1)
t_fit = time.perf_counter() - t0
e1.record(); torch.cuda.synchronize()
print('fit: host', round(1e3 * t_fit, 2), 'ms, device', round(e0.elapsed_time(e1), 2), 'ms for 20 steps')
print('device: start -> end of step 0:', round(e0.elapsed_time(events[0]), 2), 'ms; step intervals:',
      [round(events[i].elapsed_time(events[i + 1]), 2) for i in range(len(events) - 1)], '; last step -> end', round(events[-1].elapsed_time(e1), 2))
print('main/producer marks (ms after start):', [(t, round(1e3 * (a - t0), 2), round(1e3 * (b - t0), 2)) for t, a, b in marks[:14]])
print('producer: compact_item (start, end) ms after start:', [(round(1e3 * (a - t0), 1), round(1e3 * (b - t0), 1)) for a, b in prod[:6]])
print('host: enqueue times of the steps (ms after start):', [round(1e3 * (h - t0), 1) for h in hosts])
import cProfile, pstats
model._encoder_backward = orig
g3 = gen(3, 3)
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
model.fit(g3, epochs=1)
pr.disable()
pstats.Stats(pr).sort_stats('cumulative').print_stats(28)
