"""Drop-in for the reference's ``src/lib/mvnerf/model_v0.py``: ``MVVNeRFRenderer`` and ``render_view``
with the same constructor keywords, attributes and methods that ``src/train_nerf.py`` consumes,
on top of the sm_100a render path.

What is and is not here
-----------------------
* ``_call`` / ``infer`` / ``volumetric_render`` / ``train_step`` / ``render_view`` -- the hot path --
  run on the CUDA kernels (``clipnerf_b200.render.render_rays``); no CPU fallback.
* The image encoders that *produce* ``combined_features`` (``VisualFeatures``, the frozen CLIP
  SavedModel, ``CombineCLIPVisual*``) are upstream of the path and out of scope.  ``call`` /
  ``render_view`` therefore take the feature producer as a plug-in: set ``model.feature_fn`` to a
  callable ``images [(b n), H, W, 3] -> [(b n), H, W, 256]`` (any torch module), or pass
  ``combined_features`` explicitly to ``infer`` / ``_call``.
* Keras plumbing (``compile`` / ``fit`` / ``summary`` / ``store`` / ``load``) is re-stated minimally
  on torch; checkpoints are one ``torch.save`` file per reference checkpoint group
  (``<path>_<group>.pt`` instead of the TF ``.index`` / ``.data`` pair).
* Training.  All NeRF variables live in ONE ``[2, NET_PARAMS]`` parameter (``nerf_params``; the four layer modules
  hold views), and ``compile()`` with the default optimizer drives ``clipnerf_b200.trainer.RenderTrainer`` on it:
  ``train_step`` is the fused clip + warm-up + Adam kernel behind the render backward, ``fit`` uploads the generator's
  batches from pinned memory one batch ahead and replays the step from a CUDA graph (one launch per step), with the
  NCCL all-reduce of the flat gradient recorded in it when ``torch.distributed`` is initialised.  The optimizer
  groups are the reference's (``src/train_nerf.py:27-33``): the two embeddings train (Adam 1e-4 with
  ``WarmupScheduler(1e-4, 10000, 450000)``), the readouts are in no group and keep their values
  (``train_readouts=True`` changes that), the encoders belong to the plug-in (``feature_optimizer``).
  Passing a ``torch.optim.Optimizer`` or a custom loss selects a plain eager autograd step instead.
"""
import os

import numpy as np
import torch

from clipnerf_b200 import ops
from clipnerf_b200._abi import EMB_PARAMS, NET_PARAMS
from clipnerf_b200.render import render_rays

from ..data_generator.mvnerf import MVNeRFDataGenerator  # noqa: F401  (re-exported like the reference)
from ..data_generator.util import camera_parameters
from .layers import MVResNetMLPNeRFEmbedding, RenderReadout
from .nerf_utils import get_rays, optimize

_GROUPS = ('coarse_embedding', 'coarse_readout', 'fine_embedding', 'fine_readout', 'visual_features',
           'combine_clip_visual')


class _UpstreamStub(torch.nn.Module):
    """placeholder for an out-of-scope encoder group (keeps attribute names / checkpoint groups)."""

    def __init__(self, what):
        super().__init__()
        self.what = what

    # train_nerf.py:30-31 names these two sub-encoders when it builds optimizer groups
    @property
    def vision_transformer(self):
        return self

    @property
    def conv_features(self):
        return self

    def forward(self, *a, **k):
        raise NotImplementedError(f'{self.what} is upstream of the render path and not part of clipnerf_b200; '
                                  f'set model.feature_fn (see INTEGRATION.md)')

    def summary(self):
        print(f'{self.what}: out of scope (plug-in)')


class MVVNeRFRenderer(torch.nn.Module):
    def __init__(self, n_rays_train, n_rays_infer, n_views=2, n_samples=64, n_features=256,
                 embed_direction_vector=True, batch_size=1, near=0.7, far=1.5, original_image_size=(480, 640),
                 feature_fn=None, device=None, **kwargs):
        super().__init__()
        if n_samples != 64 or n_features != 256:
            raise ValueError('the sm_100a path is specialised for n_samples=64 and n_features=256 '
                             '(src/configs/nerf_model/default.yaml)')
        self.n_views = n_views
        self.coarse_embedding = MVResNetMLPNeRFEmbedding(6, 128, embed_direction_vector=embed_direction_vector,
                                                         n_views=n_views)
        self.coarse_readout = RenderReadout(4)
        self.fine_embedding = MVResNetMLPNeRFEmbedding(6, 128, embed_direction_vector=embed_direction_vector,
                                                       n_views=n_views)
        self.fine_readout = RenderReadout(4)
        # one flat buffer for every NeRF variable: row 0 = coarse net, row 1 = fine net, [embedding | readout]
        with torch.no_grad():
            init = torch.stack([torch.cat([self.coarse_embedding.flat, self.coarse_readout.flat]),
                                torch.cat([self.fine_embedding.flat, self.fine_readout.flat])])
        self.nerf_params = torch.nn.Parameter(init)
        for row, (emb, ro) in enumerate(((self.coarse_embedding, self.coarse_readout),
                                         (self.fine_embedding, self.fine_readout))):
            emb.share_storage(self, 'nerf_params', row, 0)
            ro.share_storage(self, 'nerf_params', row, EMB_PARAMS)
        self.visual_features = _UpstreamStub('VisualFeatures')
        self.clip_visual = _UpstreamStub('CLIPVisualEncoder')
        self.combine_clip_visual = _UpstreamStub('CombineCLIPVisualV0')
        self.feature_fn = feature_fn
        self.n_samples = n_samples
        self.n_rays_train = n_rays_train
        self.n_rays_infer = n_rays_infer
        self.batch_size = batch_size
        self.infer_batch_size = 1
        self.near, self.far = near, far
        self.original_image_size = tuple(original_image_size)
        self.optimizer = None
        self.loss = None
        self.loss_tracker = []
        self._trainer = None               # RenderTrainer on nerf_params (compile() with the default optimizer)
        self._pinned_ring = []             # pinned host batch buffers of fit(), allocated once
        self.max_rays_per_step = None      # larger batches are rendered in ray chunks whose gradients accumulate (eager)
        self._feature_optimizer = None
        self._generator = None
        if device is not None:
            self.to(device)

    # ------------------------------------------------------------------ helpers
    @property
    def device(self):
        return self.nerf_params.device

    def _flat(self, net):
        return self.nerf_params[0 if net == 'coarse' else 1]

    def _to_dev(self, x):
        return torch.as_tensor(np.asarray(x) if not torch.is_tensor(x) else x, dtype=torch.float32).to(self.device)

    def _rand(self, b, nr):
        if self._generator is None:
            self._generator = torch.Generator(device=self.device)
            self._generator.manual_seed(torch.initial_seed() % (2 ** 31))
        return torch.rand(b, nr, self.n_samples, device=self.device, generator=self._generator)

    # ------------------------------------------------------------------ reference API
    def encode(self, image):
        if self.feature_fn is None:
            return self.visual_features(image)
        return self.feature_fn(self._to_dev(image))

    def infer(self, inputs, batched_features, u_strat=None, u_pdf=None):
        with torch.no_grad():
            return self._call(inputs, self.n_rays_infer, self.infer_batch_size, batched_features, u_strat, u_pdf)

    def call(self, inputs, training=False, mask=None):
        src = self._to_dev(inputs[2])
        b, n = src.shape[:2]
        feats = self._features_for(src)
        return self._call(inputs, self.n_rays_train, self.batch_size, feats)

    def forward(self, inputs, training=False):
        return self.call(inputs, training=training)

    @staticmethod
    def volumetric_render(zs, density, chromacity):
        """(rgb, depth, weights) of reference :89-100 on the warp-scan compositing kernel."""
        rgbs = torch.cat([chromacity, density[..., None]], dim=-1)
        return ops.composite_fwd(zs, rgbs)

    def _call(self, inputs, n_rays, batch_size, combined_features, u_strat=None, u_pdf=None):
        """(rgb, depth, fine_rgb, fine_depth); reference :113-184.  ``n_rays`` is free (any even
        number), not the 512 the reference's input_signature pins."""
        ray_o, ray_d, images, K, Einv = [self._to_dev(t) for t in inputs]
        feats = combined_features if torch.is_tensor(combined_features) else self._to_dev(combined_features)
        b, nr = ray_o.shape[:2]
        u_strat = self._rand(b, nr) if u_strat is None else self._to_dev(u_strat)
        u_pdf = self._rand(b, nr) if u_pdf is None else self._to_dev(u_pdf)
        return render_rays(self._flat('coarse'), self._flat('fine'), (ray_o, ray_d, images, K, Einv),
                           feats.to(self.device), u_strat, u_pdf, self.near, self.far)

    # ------------------------------------------------------------------ training plumbing (Keras-like)
    def compile(self, optimizer=None, loss=None, feature_optimizer=None, train_readouts=False, target_lr=1e-4,
                warmup_steps=10000, scale_down_after=450000):
        """``optimizer`` None (default): the reference's NeRF optimizer group -- Adam on the two embeddings with
        ``WarmupScheduler(target_lr, warmup_steps, scale_down_after)`` (``train_nerf.py:22-33``), per-element gradient
        clip 1.0 (``model_v0.py:195``) -- as ``RenderTrainer`` on ``nerf_params``: fused update kernel, CUDA-graph
        ``fit``, one NCCL all-reduce per step when ``torch.distributed`` is initialised.  ``feature_optimizer``: a torch
        optimizer over the plug-in encoders' parameters (the reference's second group), stepped with the feature-map
        gradient of every step.  ``optimizer`` = a ``torch.optim.Optimizer`` or a custom ``loss``
        (``(labels, prediction) -> scalar``): plain eager autograd step on ``self.parameters()``."""
        self.loss = loss if loss is not None else (lambda y, p: torch.mean((p - y) ** 2))
        self._feature_optimizer = feature_optimizer
        if optimizer is None and loss is None:
            from clipnerf_b200.trainer import RenderTrainer
            self._trainer = RenderTrainer(self.device, params=self.nerf_params, target_lr=target_lr,
                                          warmup_steps=warmup_steps, scale_down_after=scale_down_after, clip=1.0,
                                          train_readouts=train_readouts)
            self.optimizer = self._trainer
        else:
            self._trainer = None
            self.optimizer = optimizer if optimizer is not None else torch.optim.Adam(self.parameters(), lr=target_lr,
                                                                                     eps=1e-7)

    def _features_for(self, images):
        """combined_features of a device image batch [b,nv,H,W,3] through the plug-in encoder: [b,nv,H,W,C], or
        [b,nv,H/2,W/2,C] when the plug-in stops before the final 2x bilinear UpSampling2D (layers.py:589,659) -- the
        gather then composes the upsample (CnScene.feat_up = 2)"""
        b, n = images.shape[:2]
        f = self.encode(images.reshape(b * n, *images.shape[2:]))
        return f.reshape(b, n, *f.shape[-3:])

    def _encoder_backward(self, k, feats, f_out):
        """hand the feature-map gradient of a step to the plug-in encoders (their own backward + optimizer step)"""
        if f_out is None or not f_out.requires_grad or feats.grad is None:
            return
        base = f_out._base if f_out._base is not None else f_out
        if base.grad_fn is None and base.numel() == feats.grad.numel() and base.is_contiguous():
            base.grad = feats.grad.view_as(base)     # the maps are themselves the trainable leaf: no accumulation pass
        else:
            f_out.backward(feats.grad)
        if self._feature_optimizer is not None:
            self._feature_optimizer.step()
            self._feature_optimizer.zero_grad(set_to_none=True)

    def train_step(self, data):
        """reference :186-197: loss = L(labels, rgb) + L(labels, fine_rgb); per-element clip 1.0; optimizer."""
        inputs, labels = data
        labels = self._to_dev(labels)
        if self._trainer is not None:
            dev_in = [self._to_dev(t) for t in inputs]
            f_out = self._features_for(dev_in[2])
            feats = f_out.detach().requires_grad_(f_out.requires_grad)
            b, nr = dev_in[0].shape[:2]
            loss = self._trainer.train_step(dev_in, feats, labels, self._rand(b, nr), self._rand(b, nr), self.near,
                                            self.far)
            self._encoder_backward(0, feats, f_out)
            self.loss_tracker.append(loss)
            return {'loss': loss}
        rgb, depth, fine_rgb, fine_depth = self.call(inputs, training=True)
        loss = self.loss(labels, rgb) + self.loss(labels, fine_rgb)
        variables = [p for p in self.parameters() if p.requires_grad]
        gradients = torch.autograd.grad(loss, variables, allow_unused=True)
        optimize(self.optimizer, variables, gradients, 1.0)
        self.loss_tracker.append(loss.detach())
        return {'loss': loss.detach()}

    def fit(self, generator, epochs=1, initial_epoch=0, verbose=0, graph=True):
        """Keras ``fit`` over a ``Sequence`` (``train_nerf.py:52``).  With the default optimizer every epoch runs
        ``clipnerf_b200.trainer.fit_host_batches_graph``: a background thread pulls ``generator[i]`` into pinned
        memory (Keras' enqueuer), uploads run one batch ahead on a copy stream, the step is ONE CUDA-graph launch
        (``graph=False``: eager launches) behind the uint8 -> float conversion of the uploaded source images, the
        losses come back asynchronously."""
        history = []
        for epoch in range(initial_epoch, epochs):
            if self._trainer is not None:
                from clipnerf_b200 import trainer as T
                if self._generator is None:
                    self._rand(1, 2)
                # (a generator with `compact_item` hands the source images as uint8: converted on the device)
                batches = T.PinnedBatchQueue(generator, range(len(generator)), lambda item: list(item[0]) + [item[1]],
                                             getter=getattr(generator, 'compact_item', None), ring=self._pinned_ring)
                chunked = self.max_rays_per_step is not None and self.n_rays_train > self.max_rays_per_step
                if graph and not chunked:
                    losses = T.fit_host_batches_graph(self._trainer, batches, self._features_for, self.near, self.far,
                                                      draws=self._generator, after_step=self._encoder_backward)
                else:
                    losses = T.fit_host_batches(self._trainer, batches, self._features_for, self.near, self.far,
                                                self.max_rays_per_step, draws=self._generator,
                                                after_step=self._encoder_backward)
                self.loss_tracker.extend(losses)
            else:
                losses = [float(self.train_step(generator[i])['loss']) for i in range(len(generator))]
            generator.on_epoch_end()
            history.append(float(np.mean(losses)) if len(losses) else float('nan'))
            if verbose:
                print(f'Epoch {epoch + 1}/{epochs} - loss: {history[-1]:.6f}')
        return {'loss': history}

    def summary(self):
        n = sum(p.numel() for p in self.parameters())
        print(f'MVVNeRFRenderer (clipnerf_b200): {n} trainable NeRF parameters '
              f'(coarse/fine embedding 246784 each, readouts 516 each); encoders are plug-ins')

    def _group_state(self, g):
        mod = getattr(self, g)
        return {'flat': mod.flat.detach().cpu().clone()} if hasattr(mod, 'share_storage') else mod.state_dict()

    def store(self, path):
        for g in _GROUPS:
            torch.save(self._group_state(g), f'{path}_{g}.pt')

    def load(self, path, old=False):
        """returns False unless every checkpoint group exists (reference :213-240)."""
        files = [f'{path}_{g}.pt' for g in _GROUPS]
        if not all(os.path.exists(f) for f in files):
            return False
        for g, f in zip(_GROUPS, files):
            mod, state = getattr(self, g), torch.load(f, map_location=self.device)
            if hasattr(mod, 'share_storage'):
                with torch.no_grad():
                    mod.flat.copy_(state['flat'])
            else:
                mod.load_state_dict(state)
        return True


def render_view(model, src_colors, src_camera_configs, tgt_camera_config, chunk=None):
    """full-image render of the target view (reference :243-281) -> (uint8 HxWx3, uint8 HxWx1).

    Same bookkeeping: row-major rays over the source image size, fine rgb x255 clipped and
    *truncated* to uint8, depth min-max normalised then truncated.  The reference generates the rays with
    numpy, walks the image in 600 chunks of ``n_rays_infer`` = 512 rays and copies every chunk's floats to
    the host; here the rays are generated on the device (``cn_generate_rays``), the features are encoded
    once, the image is rendered in a few large launches (``chunk`` rays each, default 65536), the uint8
    conversion runs on the device (``cn_quantize_render``) and only the two uint8 images are copied back."""
    tgt_intrinsic = np.reshape(tgt_camera_config['intrinsics'], (3, 3)).astype(np.float32)
    image_shape = src_colors[0].shape[:2]
    H, W = image_shape
    src_images = np.array([[image[..., :3] / 255.0 for image in src_colors]])
    dev = model.device
    imgs = torch.as_tensor(src_images, dtype=torch.float32, device=dev)
    feats = model.encode(imgs.reshape(-1, *imgs.shape[2:])).reshape(1, imgs.shape[1], imgs.shape[2], imgs.shape[3], -1)
    e_invs, ks = zip(*[camera_parameters(c) for c in src_camera_configs])
    K, Einv = np.array([ks], dtype=np.float32), np.array([e_invs], dtype=np.float32)
    n = H * W
    chunk = chunk or 65536
    rgbs, depths = [], []
    for i in range(0, n, chunk):
        m = min(chunk, n - i)
        m_pad = m + (m % 2)            # tiles hold two coarse rays: the padding ray is the next pixel, dropped below
        o, d = ops.generate_rays(np.asarray(tgt_camera_config['pose']), tgt_intrinsic, W, i, m_pad, dev)
        with torch.no_grad():
            _, _, fine_rgb, fine_depth = model._call((o[None], d[None], imgs, K, Einv), m_pad, 1, feats)
        rgbs.append(fine_rgb[0, :m])
        depths.append(fine_depth[0, :m])
    rgb8, depth8 = ops.quantize_render(torch.cat(rgbs), torch.cat(depths))
    return rgb8.cpu().numpy().reshape(H, W, 3), depth8.cpu().numpy().reshape(H, W, 1)
