"""Per-role timeline of CTA 0 of cn_mlp_fusion_fwd (library built with `make TRACE=1`)."""
import ctypes, sys, os
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'thesis-clip-nerf_b200'))
from clipnerf_b200 import _abi, ops
lib = _abi.load()
dev = torch.device('cuda:0')
nv = 3
tpv = 148 * 28
g = torch.Generator(device='cpu').manual_seed(0)
params = (torch.randn(_abi.NET_PARAMS, generator=g) * 0.05).to(dev)
packed = ops.pack_weights(params)
h3 = (torch.randn(nv * tpv * 128 * 128, generator=g)).to(torch.bfloat16).to(dev).view(torch.uint8)
for _ in range(3):
    ops.mlp_fusion_fwd(packed, h3, 1, nv, True)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 1024)()
lib.cn_debug_read_trace_fwd(buf, 1024)
t = np.array(buf[:], dtype=np.int64).reshape(32, 32)
names = {0: 'MMA pair start', 2: 'MMA blk0 Wa t0 issue', 3: 'MMA blk2 Wb t1 issue', 10: 'EPI tmp ready (blk0)', 11: 'EPI Eb(blk0) published',
         12: 'EPI h ready (blk0)', 13: 'EPI final start', 14: 'EPI final done', 15: 'EPI t0 pair start', 16: 'EPI t0 views landed',
         17: 'EPI t0 mean published', 19: 'EPI t0 mean loop done', 20: 'EPI t0 tmem st done'}
for it in (6, 7):
    base = t[it, 0]
    ev = sorted((int(t[it, k] - base), names[k]) for k in names if t[it, k] > 0)
    print(f'--- pair {it} (cycles relative to MMA pair start; pair period {int(t[it+1,0]-t[it,0])})')
    for c, n in ev:
        print(f'{c:8d}  {n}')
