"""fraction of texels / 2x8 work units of the feature map that no sample of the cfg2 step touches (diagnostic)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'thesis-clip-nerf_b200')]
import torch
from clipnerf_b200 import ops, synthetic
dev = torch.device('cuda:0')
nr, nv, H, W = 4096, 3, 480, 640
sc = synthetic.scene(100, nr, nv, H, W, 256, device=dev)
ro, rd, img, K, E = [t.to(dev) for t in sc['inputs']]
scene = ops.Scene(img, sc['features'], K, E)
z = ops.sample_coarse(sc['u_strat'].to(dev), 0.3, 1.3)
# fine samples are drawn near the coarse ones: use coarse + jittered copies as a proxy for the 192 samples
zz = torch.cat([z, (z + 0.004).clamp(0.3, 1.3), (z - 0.004).clamp(0.3, 1.3)], dim=-1)
pix, corner, _ = ops.project_gather_debug(scene, ro, rd, zz, gather=False)
x0, y0 = corner[..., 0].long(), corner[..., 1].long()           # [b, nv, nr, S]
occ = torch.zeros(nv, H, W, dtype=torch.bool, device=dev)
v = torch.arange(nv, device=dev).view(1, nv, 1, 1).expand_as(x0)
for dy in (0, 1):
    for dx in (0, 1):
        occ[v.reshape(-1), (y0 + dy).reshape(-1), (x0 + dx).reshape(-1)] = True
print('touched texels: %.3f' % float(occ.float().mean()))
units = occ.view(nv, H // 2, 2, W // 8, 8).any(dim=4).any(dim=2)
print('touched 2x8 units: %.3f' % float(units.float().mean()))
tiles = occ.view(nv, H // 8, 8, W // 8, 8).any(dim=4).any(dim=2)
print('touched 8x8 tiles: %.3f' % float(tiles.float().mean()))
