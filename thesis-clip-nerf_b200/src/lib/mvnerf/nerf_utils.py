"""Drop-in for the reference's ``src/lib/mvnerf/nerf_utils.py`` on torch tensors + the sm_100a C ABI.

Same function names, argument meaning and error behaviour.  Host-side ray bookkeeping stays numpy
(as in the reference); every tensor op runs a CUDA kernel of ``libclipnerf_b200.so`` -- there is no
CPU fallback.  Random draws that the reference makes inside the graph (``tf.random.uniform``) are
an optional explicit argument ``u`` so results are reproducible; when omitted they are drawn with
``torch.rand`` on the input's device.

The op-level functions here are forward-only views of the path (parity / inference / diagnostics);
training uses the fused, differentiable ``clipnerf_b200.render.render_rays`` that
``MVVNeRFRenderer._call`` calls."""
import math

import numpy as np
import torch

from clipnerf_b200 import ops


def optimize(optimizer, variables, gradients, gradients_clip=0.0):
    """per-element ``clip_by_value`` then ``apply_gradients`` (reference :8-12).  ``optimizer`` is a
    ``torch.optim.Optimizer`` over ``variables``."""
    for v, g in zip(variables, gradients):
        # a variable without a gradient this step must not see the previous step's (optimizer.step() skips None)
        v.grad = None if g is None else (g.clamp(-gradients_clip, gradients_clip) if gradients_clip > 0 else g)
    optimizer.step()


def get_rays(image_width, image_height, extrinsics, intrinsics, norm_direction_vector=True):
    """all pixels of the target view, row-major, u = column, v = row (reference :15-24)."""
    v, u = np.mgrid[0:image_height, 0:image_width].astype(np.float32)
    rays_o, rays_d = get_specific_rays(u.reshape(-1), v.reshape(-1), extrinsics, intrinsics, norm_direction_vector)
    return rays_o.reshape(image_height, image_width, 3), rays_d.reshape(image_height, image_width, 3)


def get_specific_rays(u, v, extrinsics, intrinsics, norm_direction_vector=True):
    """d = R K^-1 [u, v, 1]^T (normalised), o = camera centre (reference :27-35)."""
    pixels = np.stack((u, v, np.ones_like(u)), axis=0)
    rays_d = (extrinsics[:3, :3] @ np.linalg.inv(intrinsics[:3, :3]) @ pixels).T
    if norm_direction_vector:
        rays_d = rays_d / np.linalg.norm(rays_d, axis=1, keepdims=True)
    return np.broadcast_to(extrinsics[:3, -1], rays_d.shape), rays_d


def bbox_biased_sample(n_sample, bboxes, image_height, image_width, in_box_p=0.8):
    """(row, col) pixels: int(n*p) inside the box, the rest anywhere; numpy GLOBAL RNG, two
    ``randint`` calls in this order (reference :38-46) -- the draw sequence is part of the contract."""
    n_inside = int(n_sample * in_box_p)
    inside = np.random.randint(bboxes[:2], bboxes[2:], (n_inside, 2))
    anywhere = np.random.randint((image_height, image_width), size=(n_sample - n_inside, 2))
    return np.concatenate([inside, anywhere], axis=0)


def sample_along_ray(rays_origin, rays_direction, near, far, batch_size, n_rays, n_samples, u=None):
    """stratified samples (reference :49-61) -> (world_points [b,nr,S,3], z [b,nr,S])."""
    if u is None:
        u = torch.rand(batch_size, n_rays, n_samples, device=rays_origin.device)
    z = ops.sample_coarse(u, near, far)
    world_points = rays_origin[:, :, None, :] + z[..., None] * rays_direction[:, :, None, :]
    return world_points, z


def _scene_for(intrinsics, extr_inv, images=None, features=None, hw=(2, 2)):
    b, nv = intrinsics.shape[:2]
    dev = intrinsics.device
    if images is None:
        images = torch.zeros(b, nv, hw[0], hw[1], 3, device=dev)
    if features is None:
        features = torch.zeros(b, nv, images.shape[2], images.shape[3], 256, device=dev)
    return ops.Scene(images, features, intrinsics, extr_inv)


def compute_pixel_in_image_mv(world_points, src_intrinsics, src_extrinsics_inv):
    """(pixel_locations [b,nv,nr,np,2], camera_points_homogeneous [b,nv,nr,np,4]) (reference :64-81)."""
    b, nr, npts, _ = world_points.shape
    scene = _scene_for(src_intrinsics, src_extrinsics_inv)
    pts = world_points.reshape(b, nr * npts, 3).contiguous()
    zeros = torch.zeros_like(pts)
    # every point is a "ray" of one sample at z = 0: o + 0 * d == o exactly
    pix, _, _, cam = ops.project_gather_debug(scene, pts, zeros, torch.zeros(b, nr * npts, 1, device=pts.device),
                                              gather=False, cam=True)
    nv = scene.nv
    return pix.view(b, nv, nr, npts, 2), cam.view(b, nv, nr, npts, 4)


def world_to_camera_direction_vector_mv(world_direction_vectors, extrinsics_inverse, n_views):
    """camera-frame ray directions with homogeneous w = 1, i.e. R d + t (reference :84-105)."""
    b, nr, _ = world_direction_vectors.shape
    assert extrinsics_inverse.shape[1] == n_views
    eye = torch.eye(4, device=extrinsics_inverse.device).expand(b, n_views, 4, 4).contiguous()
    scene = _scene_for(eye, extrinsics_inverse)
    d = world_direction_vectors.contiguous()
    _, _, _, cam = ops.project_gather_debug(scene, d, torch.zeros_like(d), torch.zeros(b, nr, 1, device=d.device),
                                            gather=False, cam=True)
    return cam.view(b, n_views, nr, 4)[..., :3]


def position_encoding(position, n_freq, pos_encoding_freq):
    """[..., d] -> [..., d * n_freq * 2], per coordinate [sin f0, cos f0, sin f1, ...] (reference :108-126)."""
    return ops.position_encoding(position, n_freq, float(np.float32(pos_encoding_freq)))


def sigma_to_alpha(sigma, dists):
    """1 - exp(-dists * relu(sigma)) (reference :129-140)."""
    return 1.0 - torch.exp(-dists * torch.relu(sigma))


def sample_pdf(bins, weights, n_samples, u=None):
    """inverse-CDF draws (reference :143-176); bins [b,nr,63], weights [b,nr,62] -> [b,nr,64]."""
    if n_samples != 64 or bins.shape[-1] != 63:
        raise ValueError('the CUDA sampler is specialised for n_samples = 64 (63 bins)')
    if u is None:
        u = torch.rand(*bins.shape[:-1], n_samples, device=bins.device)
    return ops.sample_pdf(bins, weights, u)


def get_projection_features_mv(inputs, features, pixel_locations, n_rays, n_samples, batch_size):
    """bilinear gather of concat(inputs, features) at pixel_locations (reference :277-285),
    without materialising the concatenated map: -> [b, nv, nr, np, 3 + C]."""
    b, nv, h, w, _ = inputs.shape
    q = pixel_locations.reshape(b * nv, n_rays * n_samples, 2)
    g_img = ops.interpolate_bilinear(inputs.reshape(b * nv, h, w, -1), q)
    g_feat = ops.interpolate_bilinear(features.reshape(b * nv, h, w, -1), q)
    return torch.cat([g_img, g_feat], dim=-1).view(batch_size, nv, n_rays, n_samples, -1)


class WarmupScheduler:
    """linear warm-up to ``target_learning_rate``, x0.1 after ``scale_down_after`` (reference :288-300)."""

    def __init__(self, target_learning_rate, warmup_steps, scale_down_after=400000):
        self.target_learning_rate = float(target_learning_rate)
        self.warmup_steps = max(1.0, float(warmup_steps))
        self.scale_down_after = float(scale_down_after)

    def __call__(self, step):
        step = float(step)
        if step <= self.warmup_steps:
            return step / self.warmup_steps * self.target_learning_rate
        if step <= self.scale_down_after:
            return self.target_learning_rate
        return 0.1 * self.target_learning_rate


def load_pretrained_weights(path, vision_transformer):
    """The ViT encoder is upstream of the render path and not part of this package."""
    raise NotImplementedError('load_pretrained_weights: the image encoders are out of scope of clipnerf_b200; '
                              'provide combined_features from your own encoder (see INTEGRATION.md)')
