"""host cost of one batch of the end-to-end loop: MVNeRFDataGenerator.compact_item + the pinned copy (diagnostic)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]
import numpy as np, torch
from clipnerf_b200.synthetic import SyntheticNerfDataset
from src.lib.data_generator.mvnerf import MVNeRFDataGenerator
ds = SyntheticNerfDataset(n_scenes=8, n_perspectives=4, H=480, W=640, seed=1)
gen = MVNeRFDataGenerator(ds, n_rays_train=4096, batch_size=1, n_views=3, shuffle=True)
for name, fn in (('compact_item', lambda i: gen.compact_item(i)), ('__getitem__', lambda i: gen[i])):
    fn(0)
    t0 = time.perf_counter()
    for i in range(8):
        item = fn(i)
    print(name, 'ms per batch', 1e3 * (time.perf_counter() - t0) / 8)
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for i in range(8): gen.compact_item(i)
pr.disable(); pstats.Stats(pr).sort_stats('cumulative').print_stats(14)
