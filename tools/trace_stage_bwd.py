"""Per-role accounting of cn_mlp_stage_bwd at cfg2 size (fine pass: 12 288 per-view tiles, 4 096 fusion tiles) and the
CTA split between roles: cycles of every role's CTAs, cycles its producer thread spent waiting for the upstream role
(`filled`) and the downstream role (`released`), and the CUDA-event time of the launch for several splits.

    python tools/trace_stage_bwd.py [--tiles 12288] [--nv 3]
"""
import argparse
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')):
    sys.path.insert(0, p)
from clipnerf_b200 import _abi, ops, params  # noqa: E402
from clipnerf_b200._abi import ACT_TILE_BYTES, X_TILE_BYTES  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--tiles', type=int, default=12288, help='per-view stage tiles (all views)')
    ap.add_argument('--nv', type=int, default=3)
    args = ap.parse_args()
    dev = torch.device('cuda:0')
    nt, nv = args.tiles, args.nv
    tpv = nt // nv
    g = torch.Generator(device=dev).manual_seed(0)
    flat = params.glorot_init(torch.Generator().manual_seed(0), 0.02).to(dev)
    packed = ops.pack_weights(flat)

    def tiles(n, nbytes):
        return (torch.randn(n * nbytes // 2, generator=g, device=dev) * 0.5).relu_().to(torch.bfloat16).view(torch.uint8)
    acts_f, acts_u = tiles(nt, 3 * ACT_TILE_BYTES), tiles(tpv, 4 * ACT_TILE_BYTES)
    x = tiles(nt, X_TILE_BYTES)
    g_fus = (torch.randn(tpv * ACT_TILE_BYTES // 2, generator=g, device=dev) * 0.1).to(torch.bfloat16).view(torch.uint8)
    grad = torch.zeros_like(flat)
    ws = ops.stage_bwd_workspace(dev)
    lib = _abi.load()
    lib.cn_debug_stage_stats.argtypes = [ctypes.c_void_p]
    stats = torch.zeros(160, 4, dtype=torch.int64, device=dev)

    def run(stage, with_stats=False):
        lib.cn_debug_stage_stats(ctypes.c_void_p(stats.data_ptr() if with_stats else 0))
        if stage == 1:
            ops.mlp_stage_bwd(packed, 1, acts_u, 4, g_fus.clone(), grad, ws)
        else:
            ops.mlp_stage_bwd(packed, 0, acts_f, 3, g_fus, grad, ws, views=nv, b=1, scale=1.0 / nv, x_tiles=x, want_dx=True,
                              first_dense=True)
        lib.cn_debug_stage_stats(ctypes.c_void_p(0))

    def timeit(stage, reps=5):
        run(stage)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e9
        for _ in range(reps):
            if stage == 1:
                gin = g_fus.clone()
            e0.record()
            if stage == 1:
                ops.mlp_stage_bwd(packed, 1, acts_u, 4, gin, grad, ws)
            else:
                ops.mlp_stage_bwd(packed, 0, acts_f, 3, g_fus, grad, ws, views=nv, b=1, scale=1.0 / nv, x_tiles=x,
                                  want_dx=True, first_dense=True)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best

    names = ['block k+2', 'block k+1', 'block k', 'l0 dgrad', 'l0 wgrad']
    for stage, label, n in ((1, 'fusion stage', tpv), (0, 'per-view stage + first Dense', nt)):
        stats.zero_()
        run(stage, with_stats=True)
        torch.cuda.synchronize()
        s = stats.cpu()
        print(f'--- {label}: {n} tiles, {timeit(stage) * 1e3:.0f} us')
        for r in range(5):
            rows = s[(s[:, 0] == r) & (s[:, 1] > 0)]
            if len(rows) == 0:
                continue
            tiles_per_cta = n / len(rows)
            cyc = rows[:, 1].double()
            print(f'  {names[r]:10s} {len(rows):3d} CTAs  cycles mean {cyc.mean():12.0f} max {cyc.max():12.0f}  '
                  f'per tile {cyc.mean() / tiles_per_cta:8.0f}  waiting upstream {rows[:, 2].double().mean() / cyc.mean():6.1%} '
                  f'downstream {rows[:, 3].double().mean() / cyc.mean():6.1%}')
    for cost in ('5.5,3.0,2.6', '5.5,2.0,2.0', '5.5,1.5,1.5', '5.5,4.0,3.5', '5.5,2.5,1.5', '5.5,1.5,2.5'):
        os.environ['CN_STAGE_COST'] = cost
        print(f'CN_STAGE_COST={cost}: per-view stage {timeit(0) * 1e3:.0f} us')
    os.environ.pop('CN_STAGE_COST')
    # block by block, for reference
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(2):
        e0.record()
        gf = ops.mlp_block_bwd(packed, 2, acts_f, 3, 2, g_fus, grad, views=nv, scale=1.0 / nv, b=1)
        for blk in (1, 0):
            ops.mlp_block_bwd(packed, blk, acts_f, 3, blk, gf, grad, scale=1.0 / nv)
        ops.mlp_l0_wgrad(x, gf, grad, 1.0 / nv)
        ops.mlp_l0_dgrad(packed, gf, 1.0 / nv)
        e1.record()
        torch.cuda.synchronize()
    print(f'block by block: per-view stage + first Dense {e0.elapsed_time(e1) * 1e3:.0f} us')


if __name__ == '__main__':
    main()
