"""where do the tap gradients of the fused head differ from the oracle? (rows / columns of the largest errors)"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'thesis-clip-nerf_b200'))
import torch
from clipnerf_b200 import layout, ops
from oracle import nerf_oracle as O

b, n_poses, act = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
cuda = torch.device('cuda:0')
named = O.init_grasp_readout(5, bias_std=0.05)
g = torch.Generator().manual_seed(b * 1000 + n_poses)
n = n_poses * 42
n_pad = (n + 127) // 128 * 128
taps = torch.randn(4, b, n_pad, 128, generator=g) * 0.7
gy = torch.randn(b, n_poses, generator=g)
emb = [taps[s, :, :n].double().reshape(b, n_poses, 42, 128).requires_grad_(True) for s in range(4)]
p64 = {k: v.double() for k, v in named.items()}
y_ref = O.grasp_readout(emb, p64, activation=act, quant='bf16')
g_ref = torch.autograd.grad((y_ref * gy.double()).sum(), emb)
flat = torch.cat([named[f'grasp_readout.{k}'].reshape(-1) for k, _ in O.grasp_readout_param_shapes(True)]).float()
pack = ops.grasp_head_pack(flat.to(cuda))
y, st = ops.grasp_head_fwd(taps.reshape(4, b * n_pad, 128).contiguous().to(cuda), b, n_pad, n_poses, pack, act == 'elu', True)
dtap = ops.grasp_head_bwd(gy.to(cuda), st)
print('y err', float((y.cpu().double() - y_ref).abs().max() / y_ref.abs().max()))
for s in range(4):
    got = layout.tiles_to_matrix(dtap[s], 128).view(b, n_pad, 128)[:, :n].double().cpu()
    ref = g_ref[s].reshape(b, n, 128)
    d = got - ref
    rowe = d.norm(dim=-1) / ref.norm(dim=-1).max()
    worst = torch.argsort(rowe.reshape(-1), descending=True)[:12]
    print(f'tap {s}: worst rows (b*n index, pose, j, err):', [(int(w), int(w) % n // 42, int(w) % n % 42, round(float(rowe.reshape(-1)[w]), 4)) for w in worst])
    w = int(worst[0])
    dr = d.reshape(-1, 128)[w]; rr = ref.reshape(-1, 128)[w]
    print('   worst row: |d| per 16 cols', [round(float(dr[i:i + 16].norm()), 4) for i in range(0, 128, 16)], 'ref norm', float(rr.norm()))
    print('   ratio got/ref first 8 cols', [round(float(x), 3) for x in (got.reshape(-1, 128)[w][:8] / rr[:8])])
    # error by offset j and by pose
    ej = (d.reshape(b, n_poses, 42, 128).norm(dim=-1) / ref.norm(dim=-1).max()).mean(dim=(0, 1))
    print('   mean err by offset j:', [round(float(x), 4) for x in ej])
