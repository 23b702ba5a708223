"""Per-role timeline of CTA 0 of cn_mlp_feature_fwd (library built with `make TRACE=1`)."""
import ctypes, sys, os
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'thesis-clip-nerf_b200'))
from clipnerf_b200 import _abi, ops
lib = _abi.load()
dev = torch.device('cuda:0')
nt = 148 * 48
g = torch.Generator(device='cpu').manual_seed(0)
params = (torch.randn(_abi.NET_PARAMS, generator=g) * 0.05).to(dev)
packed = ops.pack_weights(params)
x = (torch.randn(nt * 128 * 384, generator=g)).to(torch.bfloat16).to(dev).view(torch.uint8)
for _ in range(3):
    ops.mlp_feature_fwd(packed, x, True)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 1024)()
lib.cn_debug_read_trace_fwd(buf, 1024)
t = np.array(buf[:], dtype=np.int64).reshape(32, 32)
names = {0: 'MMA pair start', 1: 'MMA L0 issued', 2: 'MMA blk0 Wa t0 issue', 3: 'MMA blk2 Wb t1 issue', 8: 'EPI h0 ready', 9: 'EPI Ea(blk0) published',
         10: 'EPI tmp ready (blk0)', 11: 'EPI Eb(blk0) published', 12: 'EPI h ready (blk0)', 13: 'EPI final start', 14: 'EPI final done',
         20: 'PROD first L0 load', 21: 'PROD last L0 load', 22: 'PROD last weight load'}
for it in (10, 11):
    base = t[it, 0]
    ev = sorted((int(t[it, k] - base), names[k]) for k in names if t[it, k] > 0)
    print(f'--- pair {it} (cycles relative to MMA pair start; pair period {int(t[it+1,0]-t[it,0])})')
    for c, n in ev:
        print(f'{c:8d}  {n}')
