"""Per-role timeline of CTA 0 of cn_mlp_block_bwd (library built with `make TRACE=1`)."""
import ctypes, sys, os
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'thesis-clip-nerf_b200'))
from clipnerf_b200 import _abi, ops
if os.environ.get('CN_LIB'):
    _abi.LIB_PATH = os.path.abspath(os.environ['CN_LIB'])
lib = _abi.load()
dev = torch.device('cuda:0')
nt = 148 * 24
g = torch.Generator(device='cpu').manual_seed(0)
params = (torch.randn(_abi.NET_PARAMS, generator=g) * 0.05).to(dev)
packed = ops.pack_weights(params)
acts = (torch.randn(nt * 3 * 128 * 128, generator=g).clamp_min(0)).to(torch.bfloat16).to(dev).view(torch.uint8)
gy = (torch.randn(nt * 128 * 128, generator=g) * 0.1).to(torch.bfloat16).to(dev).view(torch.uint8)
grad = torch.zeros(_abi.NET_PARAMS, device=dev)
for _ in range(3):
    ops.mlp_block_bwd(packed, 1, acts, 3, 1, gy.clone(), grad)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 1024)()
rc = lib.cn_debug_read_trace(buf, 1024)
t = np.array(buf[:], dtype=np.int64).reshape(32, 32)
names = {0: 'MMA G2 issue', 1: 'MMA G2 issued', 2: 'MMA WB issue (a2_ready seen)', 3: 'MMA WB issued+commit', 4: 'MMA T(next) done/skip',
         5: 'MMA gt_ready seen', 6: 'MMA G1 issued+commit', 7: 'MMA WA issued', 8: 'E0 start', 9: 'E0 tmem loaded', 10: 'E0 stored',
         11: 'E0 arrived', 12: 'E1 start (G2 done)', 13: 'E1 WB done seen (stores may start)', 14: 'E1 arrived', 15: 'E2 start', 17: 'E2 arrived', 18: 'EPI loop top (before the t_ready wait)', 19: 'EPI t_ready seen (before the a2buf_free wait)', 20: 'biasGy start', 21: 'biasGy end',
         22: 'biasGt start', 23: 'biasGt end', 24: 'PROD gy_staged seen', 25: 'PROD gy reload issued', 26: 'PROD free_a1 seen'}
for it in (10, 11):
    base = t[it, 8]
    ev = sorted((int(t[it, k] - base), names[k]) for k in names if t[it, k] > 0)
    print(f'--- iteration {it} (cycles relative to E0 start; tile period {int(t[it+1,8]-t[it,8])})')
    for c, n in ev:
        print(f'{c:8d}  {n}')
