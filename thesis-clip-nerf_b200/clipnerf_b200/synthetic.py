"""Synthetic scenes for benchmarks and examples: hemisphere cameras around the
workspace centre, U[0,1) images, N(0,1) feature maps, rays from a target camera.  Also a tiny
in-memory dataset with the interface ``MVNeRFDataGenerator`` expects from the reference's
``SynchronizedDatasets`` (``datasets['color'|'camera_config'].read_sample_at_idx``)."""
from __future__ import annotations

import math

import numpy as np
import torch


def polar_pose(azimuth, polar, radius, t_center=(0.6, 0.0, 0.1)):
    """camera-to-world pose on a sphere around ``t_center`` looking at it -- the construction of the
    reference's ``Affine.polar`` (dependencies/manipulation_tasks/.../transform.py:58-75)."""
    c = np.asarray(t_center, dtype=np.float64)
    t = c + radius * np.array([math.sin(polar) * math.cos(azimuth), math.sin(polar) * math.sin(azimuth),
                               math.cos(polar)])
    z_axis = (c - t) / np.linalg.norm(c - t)
    x_axis = np.cross(z_axis, np.array([0.0, 0.0, 1.0]))
    if np.linalg.norm(x_axis) == 0:
        x_axis = np.array([math.cos(azimuth), math.sin(azimuth), 0.0])
    else:
        x_axis /= np.linalg.norm(x_axis)
    y_axis = np.cross(z_axis, x_axis)
    y_axis /= np.linalg.norm(y_axis)
    pose = np.eye(4)
    pose[:3, :3] = np.stack([x_axis, y_axis, z_axis], axis=1)
    pose[:3, 3] = t
    return pose


def intrinsics(H, W, fov_deg=60.0):
    f = 0.5 * W / math.tan(math.radians(fov_deg / 2))
    return np.array([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1]], dtype=np.float64)


class _Store:
    def __init__(self, samples, n_perspectives):
        self.samples, self.n_perspectives = samples, n_perspectives

    def read_sample_at_idx(self, idx, p_idx):
        return self.samples[idx][p_idx]


class SyntheticNerfDataset:
    """``n_scenes`` scenes x ``n_perspectives`` RGBA uint8 images + camera configs, in memory."""

    def __init__(self, n_scenes=2, n_perspectives=4, H=48, W=64, seed=0):
        rs = np.random.RandomState(seed)
        K = intrinsics(H, W)
        colors, cams = [], []
        for _ in range(n_scenes):
            colors.append([rs.randint(0, 256, (H, W, 4)).astype(np.uint8) for _ in range(n_perspectives)])
            cams.append([dict(pose=polar_pose(rs.uniform(0, 2 * math.pi), rs.uniform(0.2, 1.1), rs.uniform(0.7, 0.9)),
                              intrinsics=K.reshape(-1).copy()) for _ in range(n_perspectives)])
        self.datasets = {'color': _Store(colors, n_perspectives), 'camera_config': _Store(cams, n_perspectives)}
        self._n = n_scenes

    def __len__(self):
        return self._n


def scene(seed, nr, nv, H=480, W=640, C=256, n_samples=64, device=None, pin=False):
    """one synthetic scene as torch tensors: dict(inputs=5-tuple, features, u_strat, u_pdf, target)."""
    from src.lib.data_generator.util import camera_parameters
    from src.lib.mvnerf.nerf_utils import get_specific_rays
    rs = np.random.RandomState(seed)
    g = torch.Generator().manual_seed(seed)
    K3 = intrinsics(H, W)
    poses = [polar_pose(rs.uniform(0, 2 * math.pi), rs.uniform(0.2, 1.1), rs.uniform(0.7, 0.9)) for _ in range(nv + 1)]
    n_inside = int(nr * 0.8)
    rc = np.concatenate([rs.randint((0, 0), (H, W), (n_inside, 2)), rs.randint((H, W), size=(nr - n_inside, 2))])
    o, d = get_specific_rays(rc[:, 1], rc[:, 0], poses[-1], K3.astype(np.float32))
    e_invs, ks = zip(*[camera_parameters(dict(pose=p, intrinsics=K3.reshape(-1))) for p in poses[:-1]])
    f32 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32)  # noqa: E731
    out = dict(inputs=(f32(o)[None], f32(d)[None], torch.rand((1, nv, H, W, 3), generator=g), f32(ks)[None],
                       f32(e_invs)[None]),
               features=torch.randn((1, nv, H, W, C), generator=g),
               u_strat=torch.rand((1, nr, n_samples), generator=g), u_pdf=torch.rand((1, nr, n_samples), generator=g),
               target=torch.rand((1, nr, 3), generator=g), pixels_rc=rc)
    if pin:
        out['inputs'] = tuple(t.pin_memory() for t in out['inputs'])
        for k in ('u_strat', 'u_pdf', 'target'):
            out[k] = out[k].pin_memory()
    if device is not None:
        out['features'] = out['features'].to(device)
    return out
