"""Drop-in for the reference's pose-query models: ``DeltaNGF`` (``src/lib/delta_ngf/model.py``), and through
thin subclasses ``LanguageNeRF`` (``lmvnerf/model_v4.py``) and ``GraspMVNeRF`` (``grasp_mvnerf/model.py``) --
same constructor keywords, attributes and methods that ``train_delta_ngf.py`` / ``train_language.py`` /
``train_goal.py`` consume.

``_call`` -- 42 fixed query offsets per grasp pose, projection into the source views, bilinear gather of the
259-channel map, positional encoding, fine embedding with ``complete_output``, ``GraspReadout`` -- runs on the
sm_100a kernels (``clipnerf_b200.query.query_embeddings``); no CPU fallback.  The image / CLIP encoders that
produce the feature maps are upstream plug-ins (``model.feature_fn``), exactly as for ``MVVNeRFRenderer``."""
import os

import numpy as np
import torch

from clipnerf_b200.query import pose_query_points, query_embeddings, query_grasp_success
from clipnerf_b200.trainer import allreduce_mean

from ..mvnerf.layers import MVResNetMLPNeRFEmbedding
from ..mvnerf.model_v0 import _UpstreamStub
from ..mvnerf.nerf_utils import optimize
from .layers import GraspReadout


def t_m_to_h_matrix(translations, rot_matrices):
    """[[R | t], [0 0 0 1]] (reference delta_ngf/model.py:16-25)"""
    top = torch.cat([rot_matrices, translations[..., None]], dim=-1)
    last = top.new_zeros(*top.shape[:-2], 1, 4)          # built on the device: no host copy (CUDA-graph capturable)
    last[..., 0, 3] = 1.0
    return torch.cat([top, last], dim=-2)


def quaternion_to_matrix(q):
    """tensorflow_graphics rotation_matrix_3d.from_quaternion, q = [x, y, z, w], assumed normalised"""
    x, y, z, w = q.unbind(-1)
    tx, ty, tz = 2.0 * x, 2.0 * y, 2.0 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    m = torch.stack([1.0 - (tyy + tzz), txy - twz, txz + twy, txy + twz, 1.0 - (txx + tzz), tyz - twx,
                     txz - twy, tyz + twx, 1.0 - (txx + tyy)], dim=-1)
    return m.reshape(*q.shape[:-1], 3, 3)


def t_q_to_h_matrix(translations, quaternions):
    return t_m_to_h_matrix(translations, quaternion_to_matrix(quaternions))


def pose_matrices(translations, rotations, rotation_representation):
    """``compute_matrices`` (reference :172-183); the 6-d form does NOT orthogonalise the second vector"""
    if rotation_representation == 'quaternion':
        return t_q_to_h_matrix(translations, rotations)
    if rotation_representation == '6d':
        r_1 = rotations[..., :3] / rotations[..., :3].norm(dim=-1, keepdim=True)
        r_2 = rotations[..., 3:] / rotations[..., 3:].norm(dim=-1, keepdim=True)
        r_3 = torch.linalg.cross(r_1, r_2, dim=-1)
        return t_m_to_h_matrix(translations, torch.stack([r_1, r_2, r_3], dim=-1))
    raise ValueError('Unknown rotation representation: ' + rotation_representation)


def build_transforms_to_check(n_5d_poses=7):
    """6 gripper bases x n_5d_poses steps of 5 mm along the local z axis (reference :63-97), [6*n, 4, 4] fp32.
    Euler 'xyz' rotations through scipy, like the reference's ``manipulation_tasks.transform.Affine``."""
    from scipy.spatial.transform import Rotation

    def affine(t, euler=None):
        m = np.eye(4)
        m[:3, 3] = t
        if euler is not None:
            m[:3, :3] = Rotation.from_euler('xyz', euler).as_matrix()
        return m
    ox, oy, oz = 0.02, 0.015, 0.0125
    step_size = (ox - 0.005) / ((n_5d_poses - 1) / 2)
    bases = [affine([0, oy, 0]), affine([0, -oy, 0]),
             affine([-ox, oy, oz], [0.0, np.pi / 2, 0.0]), affine([ox, oy, oz], [0.0, -np.pi / 2, 0.0]),
             affine([-ox, -oy, oz], [0.0, np.pi / 2, 0.0]), affine([ox, -oy, oz], [0.0, -np.pi / 2, 0.0])]
    c = int((n_5d_poses - 1) / 2)
    steps = [affine([0.0, 0.0, i * step_size]) for i in range(-c, c + 1)]
    return torch.tensor(np.array([b @ t for b in bases for t in steps]), dtype=torch.float32)


def cosine_similarity(y_true, y_pred):
    """``tf.keras.losses.CosineSimilarity(axis=-1)``"""
    def l2n(x):
        return x * torch.rsqrt(torch.clamp((x * x).sum(-1, keepdim=True), min=1e-12))
    return -(l2n(y_true) * l2n(y_pred)).sum(-1).mean()


class _PoseQueryModel(torch.nn.Module):
    """what DeltaNGF / LanguageNeRF / GraspMVNeRF share: the fine embedding, the affordance head, the query
    offsets, ``_call`` and the checkpoint groups."""
    _groups = ('fine_embedding', 'visual_features', 'grasp_readout')
    _backbone_groups = ('fine_embedding', 'visual_features')

    def _init_common(self, n_views, n_5d_poses, readout, tf32_head=True):
        readout.tf32 = bool(tf32_head)        # see GraspReadout: TensorFloat-32 matmuls, as TensorFlow runs them
        self.n_views = n_views
        self.fine_embedding = MVResNetMLPNeRFEmbedding(6, 128, n_views=n_views, embed_direction_vector=True,
                                                       complete_output=True)
        self.visual_features = _UpstreamStub('VisualFeatures')
        self.grasp_readout = readout
        self.n_5d_poses = n_5d_poses
        self.register_buffer('transforms_to_check', build_transforms_to_check(n_5d_poses), persistent=False)
        self.n_transforms_to_check = self.transforms_to_check.shape[0]
        self.feature_fn = None
        self.optimizer = None
        self.loss = None
        self.fused_head = True                # cn_grasp_head_* whenever no parameter / feature-map gradient is asked for

    @property
    def device(self):
        return self.fine_embedding.flat.device

    def _to_dev(self, x):
        return torch.as_tensor(np.asarray(x) if not torch.is_tensor(x) else x, dtype=torch.float32).to(self.device)

    def _net_params(self):
        # flat layout of the kernels: embedding followed by a RenderReadout tail the query path never reads
        return torch.cat([self.fine_embedding.flat, self.fine_embedding.flat.new_zeros(516)])

    def encode(self, src_images):
        """images [b, n, H, W, 3] -> feature maps [b, n, H, W, 256] through the plug-in encoder"""
        b, n = src_images.shape[:2]
        flat = src_images.reshape(b * n, *src_images.shape[2:])
        feats = self.visual_features(flat) if self.feature_fn is None else self.feature_fn(flat)
        return feats.reshape(b, n, *feats.shape[1:])

    def _query(self, src_images, src_intrinsics, src_extrinsics_inv, transforms, n_points, batched_features):
        """reference ``_call`` body (delta_ngf/model.py:188-245): grasp_success [b, n_points]"""
        images, K, Einv = self._to_dev(src_images), self._to_dev(src_intrinsics), self._to_dev(src_extrinsics_inv)
        feats = batched_features if torch.is_tensor(batched_features) else self._to_dev(batched_features)
        transforms = transforms if torch.is_tensor(transforms) else self._to_dev(transforms)
        b = images.shape[0]
        if transforms.shape[0] != b or transforms.shape[1] != n_points:
            raise ValueError(f'transforms must be [{b}, {n_points}, 4, 4], got {tuple(transforms.shape)}')
        pts, dirs = pose_query_points(transforms.to(self.device), self.transforms_to_check)
        head = self.grasp_readout
        trains = torch.is_grad_enabled() and (any(p.requires_grad for p in head.parameters())
                                              or self.fine_embedding.flat.requires_grad or feats.requires_grad)
        if self.fused_head and head.fused_available(self.n_transforms_to_check, trains):
            # inference / pose optimisation (frozen field): embedding AND head on the CUDA kernels, gradient to the poses
            return query_grasp_success(self._net_params(), head.packed(), images, K, Einv, feats.to(self.device), pts, dirs,
                                       n_points, elu=head.activation == 'elu')
        emb = query_embeddings(self._net_params(), images, K, Einv, feats.to(self.device), pts, dirs)
        emb = [e.reshape(b, n_points, self.n_transforms_to_check, 128) for e in emb]
        return self.grasp_readout(emb)

    # ---- checkpoints: one torch.save file per reference checkpoint group (<path>_<group>.pt)
    def load_backbone(self, path, verbose=True):
        for g in self._backbone_groups:
            if not os.path.exists(f'{path}_{g}.pt'):
                if verbose:
                    print(f'{path}_{g}.pt does not exist')
                return False
        for g in self._backbone_groups:
            getattr(self, g).load_state_dict(torch.load(f'{path}_{g}.pt', map_location=self.device))
        return True

    def store(self, path):
        for g in self._groups:
            torch.save(getattr(self, g).state_dict(), f'{path}_{g}.pt')

    def load(self, path):
        if self.load_backbone(path, verbose=False):
            f = f'{path}_grasp_readout.pt'
            if not os.path.exists(f):
                print(f'{f} does not exist')
                return False
            self.grasp_readout.load_state_dict(torch.load(f, map_location=self.device))
            return True
        print(f'backbone models at {path} do not exist')
        return False

    def compile(self, optimizer=None, loss=None):
        """only the affordance head trains (reference train_step: ``tape.watch(grasp_readout.trainable_variables)``)"""
        # capturable on a CUDA device: the step can be recorded into a CUDA graph (capture_train_step)
        self.optimizer = optimizer if optimizer is not None else torch.optim.Adam(
            self.grasp_readout.parameters(), lr=1e-4, eps=1e-7, capturable=self.device.type == 'cuda')
        self.loss = loss if loss is not None else (lambda y, p: torch.mean((p - y) ** 2))

    def fit(self, generator, epochs=1, initial_epoch=0, verbose=0, graph=False):
        """Keras ``fit`` over a ``Sequence``.  ``graph=True`` (models with ``capture_train_step``): the first batch's
        step is recorded into a CUDA graph, every batch is copied into its static buffers and replayed -- the ~700 small
        launches of a step (double backward of the head, tangent pass) then cost the host one launch."""
        history = []
        replay = None
        for epoch in range(initial_epoch, epochs):
            if graph and hasattr(self, 'capture_train_step') and len(generator):
                logs, first = [], 0
                if replay is None:
                    # the very first step runs eagerly (it creates the optimizer state), the recording follows it
                    # without training: every batch is trained on exactly once
                    logs.append(self.train_step(generator[0]))
                    replay = self.capture_train_step(generator[0], warmup=0)
                    first = 1
                logs += [{k: v.clone() for k, v in replay(generator[i]).items()} for i in range(first, len(generator))]
            else:
                logs = [self.train_step(generator[i]) for i in range(len(generator))]
            if hasattr(generator, 'on_epoch_end'):
                generator.on_epoch_end()
            history.append({k: float(np.mean([float(l[k]) for l in logs])) for k in (logs[0] if logs else {})})
            if verbose:
                print(f'Epoch {epoch + 1}/{epochs} - {history[-1]}')
        return history

    def summary(self):
        n = sum(p.numel() for p in self.grasp_readout.parameters())
        print(f'{type(self).__name__} (clipnerf_b200): fine embedding 246784 (frozen), GraspReadout {n} trainable; '
              f'encoders are plug-ins')


class DeltaNGF(_PoseQueryModel):
    def __init__(self, n_features=256, original_image_size=(480, 640), n_points_train=5, n_views=1, n_5d_poses=7,
                 pretrained=True, batch_size=1, fixed_orientation=None, rotation_representation='quaternion',
                 softmax_before_loss=False, device=None, tf32_head=True, **kwargs):
        super().__init__()
        if not pretrained:
            raise NotImplementedError('GraspMVNeRF only supports pretrained models')
        if n_features != 256:
            raise ValueError('the sm_100a path is specialised for n_features=256')
        if rotation_representation not in ('quaternion', '6d'):
            raise ValueError('Unknown rotation representation: ' + rotation_representation)
        self._init_common(n_views, n_5d_poses, GraspReadout(use_bias=True), tf32_head)
        self.pretrained = pretrained
        self.n_points_train = n_points_train
        self.batch_size = batch_size
        self.softmax_before_loss = softmax_before_loss
        self.rotation_representation = rotation_representation
        self.original_image_size = tuple(original_image_size)
        self.translations = torch.nn.Parameter(torch.zeros(batch_size, n_points_train, 3))
        r_dim = 4 if rotation_representation == 'quaternion' else 6
        self.rotations = torch.nn.Parameter(torch.zeros(batch_size, n_points_train, r_dim))
        self.pose_variables = [self.translations, self.rotations]
        self.mse = lambda y, p: torch.mean((p - y) ** 2)
        self.cosine_similarity = cosine_similarity
        self.metrics_log = []
        if device is not None:
            self.to(device)

    # the reference names the rotation variable after its representation
    @property
    def quaternions(self):
        return self.rotations

    @property
    def six_d_rotations(self):
        return self.rotations

    def compute_matrices(self):
        return pose_matrices(self.translations, self.rotations, self.rotation_representation)

    def set_pose(self, translations, rotations):
        with torch.no_grad():
            self.translations.copy_(self._to_dev(translations))
            self.rotations.copy_(self._to_dev(rotations))

    def call(self, inputs, training=False, mask=None):
        feats = self.encode(self._to_dev(inputs[4]))
        return self._call(inputs, self.compute_matrices(), self.n_points_train, feats)

    def forward(self, inputs, training=False):
        return self.call(inputs, training=training)

    def infer(self, inputs, transforms, n_points_infer, batched_features):
        return self._call(inputs, transforms, n_points_infer, batched_features)

    def _call(self, inputs, transforms, n_points, batched_features):
        return self._query(inputs[4], inputs[5], inputs[6], transforms, n_points, batched_features)

    def train_step(self, data):
        """reference :247-309: landscape loss on the predicted grasp success at the landscape poses (inputs[0:2])
        plus the gradient-matching terms -- cosine similarity between d prediction / d pose at the trajectory
        poses (inputs[2:4]) and the labelled directions -- all differentiated w.r.t. the affordance head only.
        The pose gradient comes from the CUDA backward; its derivative w.r.t. the head's weights goes through
        the tangent pass of the query call (``clipnerf_b200.query_tangent``)."""
        inputs, labels = data
        readout_vars = list(self.grasp_readout.parameters())
        feats = self.encode(self._to_dev(inputs[4])).detach()
        losses = self.compute_losses(inputs, labels, feats)
        gradients = torch.autograd.grad(losses['loss'], readout_vars, allow_unused=True)
        # data parallel over scenes / poses: every rank steps on the mean gradient (one collective, no-op on one rank)
        gradients = allreduce_mean(gradients, readout_vars, getattr(self, 'process_group', None))
        optimize(self.optimizer, readout_vars, gradients, 1.0)
        out = {'landscape_loss': losses['landscape_loss'].detach(), 'grad_loss_t': losses['grad_loss_t'].detach(),
               'grad_loss_r': losses['grad_loss_r'].detach(), 'pred': losses['prediction'].detach().mean()}
        self.metrics_log.append({k: float(v) for k, v in out.items()})
        return out

    def capture_train_step(self, data, warmup=2):
        """CUDA-graph form of ``train_step``: both ``_call``s, the pose gradient with its double backward through the
        head and the tangent pass, the (data-parallel) gradient mean, clip and the optimizer step are recorded once and
        replayed.  ``data`` fixes the shapes; the optimizer must be capturable (``compile()``'s default on CUDA) and the
        feature plug-in must not synchronise with the host.  Returns ``replay(data=None) -> metrics`` (device tensors,
        overwritten by the next replay); ``warmup`` >= 1 eager steps on ``data`` precede the capture and DO train (the
        optimizer state has to exist before the capture: a state tensor created while recording would be re-initialised
        by every replay)."""
        has_state = len(getattr(self.optimizer, 'state', {})) > 0
        assert warmup >= 1 or has_state, 'the optimizer state must be created by an eager step before the capture'
        inputs, labels = data
        s_in = [self._to_dev(x).clone() for x in inputs[:7]]
        s_lab = [self._to_dev(x).clone() for x in labels]
        readout_vars = list(self.grasp_readout.parameters())

        def step():
            feats = self.encode(s_in[4]).detach()
            losses = self.compute_losses(s_in, s_lab, feats)
            gradients = torch.autograd.grad(losses['loss'], readout_vars, allow_unused=True)
            gradients = allreduce_mean(gradients, readout_vars, getattr(self, 'process_group', None))
            optimize(self.optimizer, readout_vars, gradients, 1.0)
            return {'landscape_loss': losses['landscape_loss'].detach(), 'grad_loss_t': losses['grad_loss_t'].detach(),
                    'grad_loss_r': losses['grad_loss_r'].detach(), 'pred': losses['prediction'].detach().mean()}
        from clipnerf_b200 import layout
        layout._elem_index_on(128, str(self.device))       # index table of the tangent pass: on the device before capture
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                step()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = step()

        def replay(data=None):
            if data is not None:
                for dst, src in zip(s_in + s_lab, list(data[0][:7]) + list(data[1])):
                    dst.copy_(self._to_dev(src), non_blocking=True)
            graph.replay()
            return out
        replay.graph = graph
        return replay

    def compute_losses(self, inputs, labels, feats):
        """the loss terms of ``train_step`` (reference :258-291) as differentiable tensors"""
        self.set_pose(inputs[0], inputs[1])
        y_pred = self._call(inputs, self.compute_matrices().detach(), self.n_points_train, feats)
        if self.softmax_before_loss:
            y_pred = torch.softmax(y_pred, dim=-1)
        landscape_loss = self.loss(self._to_dev(labels[0]), y_pred)
        self.set_pose(inputs[2], inputs[3])
        prediction = self._call(inputs, self.compute_matrices(), self.n_points_train, feats)
        output_gradients = torch.autograd.grad(prediction.sum(), self.pose_variables, create_graph=True)
        loss_t = self.cosine_similarity(self._to_dev(labels[1]), output_gradients[0])
        lab_r = self._to_dev(labels[2])
        if self.rotation_representation == 'quaternion':
            loss_r = self.cosine_similarity(lab_r, output_gradients[1])
        else:
            loss_r = self.cosine_similarity(lab_r[..., :3], output_gradients[1][..., :3]) + \
                self.cosine_similarity(lab_r[..., 3:], output_gradients[1][..., 3:])
        return {'loss': loss_t + loss_r + landscape_loss, 'landscape_loss': landscape_loss, 'grad_loss_t': loss_t,
                'grad_loss_r': loss_r, 'prediction': prediction, 'output_gradients': output_gradients}
