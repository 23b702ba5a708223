"""Per-role timeline of CTA 0 of cn_mlp_chain_fwd (library built with `make TRACE=1`): the render / inference form
(nothing saved) and the training form (7 activation tiles per tile saved)."""
import ctypes, sys, os
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'thesis-clip-nerf_b200'))
from clipnerf_b200 import _abi, ops
if os.environ.get('CN_LIB'):
    _abi.LIB_PATH = os.path.abspath(os.environ['CN_LIB'])
lib = _abi.load()
dev = torch.device('cuda:0')
nt = 148 * 48
g = torch.Generator(device='cpu').manual_seed(0)
params = (torch.randn(_abi.NET_PARAMS, generator=g) * 0.05).to(dev)
packed = ops.pack_weights(params)
x = (torch.randn(nt * 128 * 384, generator=g)).to(torch.bfloat16).to(dev).view(torch.uint8)
names = {0: 'MMA pair start', 1: 'MMA L0 issued', 2: 'MMA blk0 Wa t0 issue', 3: 'MMA last blk Wb t1 issue', 8: 'EPI h0 ready',
         9: 'EPI Ea(blk0) published', 10: 'EPI tmp ready (blk0)', 11: 'EPI Eb(blk0) published', 12: 'EPI h ready (blk0)',
         4: '  blk1 t0: MMA woke on act_ready (Wa)', 5: '  blk1 t0: MMA Wa issued + committed', 6: '  blk1 t0: EPI woke on tmp_ready',
         28: '  blk1 t0: EPI E_b stores done', 29: '  blk1 t0: EPI E_b published', 30: '  blk1 t0: MMA woke on act_ready (Wb)',
         31: '  blk1 t0: MMA Wb issued + committed', 7: '  blk1 t0: EPI woke on h_ready', 13: 'EPI per-view blocks done', 18: 'EPI (tile 1) per-view blocks done', 23: 'EPI (tile 1) readout done', 14: 'EPI readout done', 20: 'XPROD first X load', 21: 'PROD last L0 weight load',
         22: 'PROD last weight load'}
for save in (False, True):
    for _ in range(3):
        ops.mlp_chain_fwd(packed, x, save)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.mlp_chain_fwd(packed, x, save)
    e1.record()
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * 1024)()
    lib.cn_debug_read_trace_fwd(buf, 1024)
    t = np.array(buf[:], dtype=np.int64).reshape(32, 32)
    print(f'===== save_acts={save}: {e0.elapsed_time(e1) * 1e3:.0f} us for {nt} tiles ({nt // 148 // 2} pairs per CTA)')
    for it in (10, 11):
        base = t[it, 0]
        ev = sorted((int(t[it, k] - base), names[k]) for k in names if t[it, k] > 0)
        print(f'--- pair {it} (cycles relative to MMA pair start; pair period {int(t[it + 1, 0] - t[it, 0])})')
        for c, n in ev:
            print(f'{c:8d}  {n}')
        if t[it, 31] != 0 and False: pass
        print(f'          MMA thread in the first Dense (CN_TRACE_FINE): {int(t[it, 30])} cycles waiting for weight slabs, {int(t[it, 31])} for X slabs')
        print(f'          traced epilogue thread (tile 0): {int(t[it, 24])} cycles in the block epilogues (of which {int(t[it, 27])} in publish = fence + arrive), '
              f'{int(t[it, 25])} waiting for tmp (6 first Denses), {int(t[it, 26])} waiting for h (6 second Denses)')
