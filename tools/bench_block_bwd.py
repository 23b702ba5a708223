"""Device time of cn_mlp_block_bwd alone (CUDA events) at the cfg2 fine-pass sizes: per-view blocks (12 288 tiles) and fusion
blocks (4 096 tiles), with the weight gradients.  CN_LIB selects another build of the library (before / after)."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'thesis-clip-nerf_b200'))
from clipnerf_b200 import _abi, ops
if os.environ.get('CN_LIB'):
    _abi.LIB_PATH = os.path.abspath(os.environ['CN_LIB'])
dev = torch.device('cuda:0')
g = torch.Generator().manual_seed(0)
packed = ops.pack_weights((torch.randn(_abi.NET_PARAMS, generator=g) * 0.05).to(dev))
sms = torch.cuda.get_device_properties(0).multi_processor_count
for name, nt in (('per-view blocks, cfg2 fine', 12288), ('fusion blocks, cfg2 fine', 4096)):
    acts = torch.randn(nt * 3 * 128 * 128, generator=g).clamp_min(0).to(torch.bfloat16).to(dev).view(torch.uint8)
    gy = (torch.randn(nt * 128 * 128, generator=g) * 0.1).to(torch.bfloat16).to(dev).view(torch.uint8)
    grad = torch.zeros(_abi.NET_PARAMS, device=dev)
    ts = []
    for i in range(8):
        gin = gy.clone()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ops.mlp_block_bwd(packed, 1, acts, 3, 1, gin, grad); e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            ts.append(e0.elapsed_time(e1) * 1e3)
    us = sorted(ts)[len(ts) // 2]
    tflops = nt * 128 * 4 * 2 * 128 * 128 / (us * 1e-6) / 1e12
    print(f'cn_mlp_block_bwd  {name:28s} {us:8.1f} us  {nt:6d} tiles  {us * 1965.0 / (nt / sms):7.0f} cycles per tile  {tflops:6.1f} TFLOP/s credited')
    del acts, gy
