"""Device time of the forward MLP kernels alone (CUDA events, L2 flushed between launches by the size of the inputs):
per-view stage and fusion stage at the cfg2 fine-pass size, chained kernel at the cfg3(i) fine-pass size (saving) and
at a render_view chunk (nothing saved).  One line per kernel: us per launch, cycles per tile pair at the SM clock."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'thesis-clip-nerf_b200'))
from clipnerf_b200 import _abi, ops
if os.environ.get('CN_LIB'):            # another build of the library (before / after comparisons)
    _abi.LIB_PATH = os.path.abspath(os.environ['CN_LIB'])

dev = torch.device('cuda:0')
g = torch.Generator().manual_seed(0)
packed = ops.pack_weights((torch.randn(_abi.NET_PARAMS, generator=g) * 0.05).to(dev))
sms = torch.cuda.get_device_properties(0).multi_processor_count


def x_tiles(nt):
    return (torch.randn(nt * 128 * 384, generator=g) * 0.5).to(torch.bfloat16).to(dev).view(torch.uint8)


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts)[len(ts) // 2]


def report(name, us, nt, mhz=1965.0):
    pairs = (nt / sms) / 2
    print(f'{name:58s} {us:8.1f} us  {nt:6d} tiles  {us * mhz / pairs:8.0f} cycles per pair (at {mhz:.0f} MHz)')


nt = 4096 * 128 * 3 // 128
x = x_tiles(nt)
report('cn_mlp_feature_fwd  cfg2 fine, saving 3 tiles', timed(lambda: ops.mlp_feature_fwd(packed, x, True)), nt)
report('cn_mlp_feature_fwd  cfg2 fine, nothing saved', timed(lambda: ops.mlp_feature_fwd(packed, x, False)), nt)
h3, _ = ops.mlp_feature_fwd(packed, x, False)
report('cn_mlp_fusion_fwd   cfg2 fine (nv = 3), saving 4 tiles', timed(lambda: ops.mlp_fusion_fwd(packed, h3, 1, 3, True)), nt // 3)
report('cn_mlp_fusion_fwd   cfg2 fine (nv = 3), nothing saved', timed(lambda: ops.mlp_fusion_fwd(packed, h3, 1, 3, False)), nt // 3)
del x, h3
nt = 8192
x = x_tiles(nt)
report('cn_mlp_chain_fwd    cfg3(i) fine, saving 7 tiles', timed(lambda: ops.mlp_chain_fwd(packed, x, True)), nt)
del x
nt = 65536 * 128 // 128
x = x_tiles(nt)
report('cn_mlp_chain_fwd    render_view chunk (65536 x 128), nothing saved', timed(lambda: ops.mlp_chain_fwd(packed, x, False)), nt)
