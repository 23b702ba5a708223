"""Drop-in for the reference's ``src/lib/delta_ngf/grasp_optimizer.py`` (= ``lmvnerf/grasp_optimizer.py``):
``DNGFOptimizer`` -- test-time gradient ascent of grasp poses on the learned field.

Every ``optimize_pose`` step evaluates ``n_initial_guesses`` poses x 42 query offsets in every source image
through the CUDA query path (gather + tcgen05 MLP forward, then the pose-gradient backward), sums the grasp
success over the image batch (reference :172-175) and applies the two optimizers (translation / rotation)
to the pose variables with per-element clipping at 1.0, followed by ``post_process`` (workspace clip,
re-normalisation of the rotation variables)."""
import numpy as np
import torch

from ..mvnerf.nerf_utils import optimize
from .model import pose_matrices, t_m_to_h_matrix, t_q_to_h_matrix  # noqa: F401  (re-exported like the reference)


def _random_poses(workspace_bounds, n, rng):
    """``Affine.random(workspace_bounds)`` (manipulation_tasks/transform.py:31-55): translation uniform in the
    bounds, rotation from uniform 'xyz' Euler angles in [0, 2 pi) -> (translations [n,3], quaternions xyzw [n,4],
    rotation matrices [n,3,3])"""
    from scipy.spatial.transform import Rotation
    t_b = np.array(workspace_bounds)
    ts = rng.uniform(t_b[:, 0], t_b[:, 1], size=(n, 3))
    rot = Rotation.from_euler('xyz', rng.uniform(0, 2 * np.pi, size=(n, 3)))
    return ts, rot.as_quat(), rot.as_matrix()


class DNGFOptimizer(torch.nn.Module):
    def __init__(self, nerf_grasper, workspace_bounds, n_initial_guesses=32, n_images=3, fixed_orientation=None,
                 clip_translation=False, rotation_representation='quaternion', **kwargs):
        super().__init__()
        if rotation_representation not in ('quaternion', '6d'):
            raise ValueError('Unknown rotation representation: ' + rotation_representation)
        self.workspace_bounds = np.array(workspace_bounds)
        self.nerf_grasper = nerf_grasper
        # test-time optimisation differentiates the pose variables only (reference :159-178: the tape's sources
        # are self.pose_variables); the trained field is frozen so that no weight gradient is computed
        for p in nerf_grasper.parameters():
            p.requires_grad_(False)
        self.n_initial_guesses = n_initial_guesses
        self.n_images = n_images
        batch_size = self.n_images / self.nerf_grasper.n_views
        assert (batch_size % 1) == 0
        self.batch_size = int(batch_size)
        dev = nerf_grasper.device
        self.translations = torch.nn.Parameter(torch.zeros(1, n_initial_guesses, 3, device=dev))
        self.rotation_representation = rotation_representation
        r_dim = 4 if rotation_representation == 'quaternion' else 6
        self.rotations = torch.nn.Parameter(torch.zeros(1, n_initial_guesses, r_dim, device=dev))
        self.pose_variables = [[self.translations], [self.rotations]]
        self.clip_translation = clip_translation
        self.optimizer = None
        self.loss_tracker = []
        self._rng = np.random.default_rng(0)
        self._bounds = None

    @property
    def quaternions(self):
        return self.rotations

    @property
    def six_d_rotations(self):
        return self.rotations

    def compile(self, optimizer=None):
        """``optimizer``: [translation optimizer, rotation optimizer] over ``pose_variables[0 / 1]``"""
        if optimizer is None:
            cap = self.translations.is_cuda          # capturable: the step can be recorded into a CUDA graph (capture())
            optimizer = [torch.optim.Adam(self.pose_variables[0], lr=1e-3, capturable=cap),
                         torch.optim.Adam(self.pose_variables[1], lr=1e-2, capturable=cap)]
        assert len(optimizer) == 2
        self.optimizer = optimizer

    def set_initial_guesses(self, initial_guesses):
        assert len(initial_guesses) == 2
        assert tuple(initial_guesses[0].shape) == (1, self.n_initial_guesses, 3)
        assert tuple(initial_guesses[1].shape) == tuple(self.rotations.shape)
        with torch.no_grad():
            self.translations.copy_(torch.as_tensor(np.asarray(initial_guesses[0]), dtype=torch.float32))
            self.rotations.copy_(torch.as_tensor(np.asarray(initial_guesses[1]), dtype=torch.float32))

    def generate_initial_guesses(self, workspace_bounds=None, n_initial_guesses=None, batch_size=1):
        workspace_bounds = self.workspace_bounds if workspace_bounds is None else workspace_bounds
        n = self.n_initial_guesses if n_initial_guesses is None else n_initial_guesses
        ts, rs = [], []
        for _ in range(batch_size):
            t, quat, mat = _random_poses(workspace_bounds, n, self._rng)
            ts.append(t)
            rs.append(quat if self.rotation_representation == 'quaternion'
                      else np.concatenate([mat[:, :, 0], mat[:, :, 1]], axis=-1))
        return [np.array(ts), np.array(rs)]

    def compute_matrices(self):
        return pose_matrices(self.translations, self.rotations, self.rotation_representation)

    def get_results(self):
        """the optimised poses as 4x4 numpy matrices (the reference wraps each in ``Affine.from_matrix``)"""
        return [m.detach().cpu().numpy() for m in self.compute_matrices()[0]]

    def call(self, inputs, training=False, mask=None):
        """inputs = [(images, K, E^-1, clip_tokens), features] (reference :104-111)"""
        matrices = self.compute_matrices().repeat(self.batch_size, 1, 1, 1)
        src = inputs[0]
        return self.nerf_grasper.infer((None, None, None, None, src[0], src[1], src[2], src[3] if len(src) > 3 else None),
                                       matrices, self.n_initial_guesses, inputs[1])

    def forward(self, inputs, training=False):
        return self.call(inputs, training=training)

    def _regroup(self, inputs, features):
        """reference :160-171: with several image batches the view axis becomes the batch axis"""
        if self.batch_size != 1:
            r_inputs = [torch.as_tensor(np.asarray(t) if not torch.is_tensor(t) else t).transpose(0, 1) for t in inputs[:3]]
            r_inputs.append(inputs[3] if len(inputs) > 3 else None)
            features = features.transpose(0, 1)
            return r_inputs, features
        return inputs, features

    def compute_current_grasp_success(self, inputs, features):
        r_inputs, r_features = self._regroup(inputs, features)
        with torch.no_grad():
            grasp_success = self([r_inputs, r_features])
        return grasp_success.sum(dim=0, keepdim=True).transpose(0, 1)          # 'b np -> np b'

    def post_process(self):
        with torch.no_grad():
            if self.clip_translation:
                if self._bounds is None or self._bounds[0].device != self.translations.device:
                    dev = self.translations.device
                    self._bounds = (torch.as_tensor(self.workspace_bounds[:, 0], dtype=torch.float32, device=dev),
                                    torch.as_tensor(self.workspace_bounds[:, 1], dtype=torch.float32, device=dev))
                lo, hi = self._bounds
                self.translations.copy_(torch.minimum(torch.maximum(self.translations, lo), hi))
            r = self.rotations
            if self.rotation_representation == 'quaternion':
                r.copy_(r / r.norm(dim=-1, keepdim=True))
            else:
                r.copy_(torch.cat([r[..., :3] / r[..., :3].norm(dim=-1, keepdim=True),
                                   r[..., 3:] / r[..., 3:].norm(dim=-1, keepdim=True)], dim=-1))

    def _step(self, r_inputs, r_features, train_config):
        grasp_success = self([r_inputs, r_features])
        grasp_success = grasp_success.sum(dim=0, keepdim=True).transpose(0, 1)
        loss = -grasp_success
        grads = torch.autograd.grad(loss.sum(), [self.translations, self.rotations])
        for i, t_c in enumerate(train_config):
            if t_c:
                optimize(self.optimizer[i], self.pose_variables[i], [grads[i]], 1.0)
        self.post_process()
        return grasp_success

    def capture(self, inputs, features, train_config=(True, True), warmup=3):
        """CUDA-graph form of ``optimize_pose`` for the test-time loop (32+ steps on fixed images, reference
        src/utils/optimization.py:136-153): one step -- ~20 kernels of the query path, ~60 torch ops of the head,
        the two optimizers and ``post_process`` -- is captured once and replayed; ``inputs`` / ``features`` must be
        device tensors that stay alive, the optimizers must be capturable (``torch.optim.Adam(..., capturable=True)``).
        Returns ``replay() -> {'loss': grasp_success [np, 1]}`` (the tensor is overwritten by the next replay).
        The ``warmup`` eager steps that precede the capture DO update the poses."""
        r_inputs, r_features = self._regroup(inputs, features)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._step(r_inputs, r_features, train_config)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = self._step(r_inputs, r_features, train_config).detach()

        def replay():
            graph.replay()
            return {'loss': out}
        replay.graph = graph
        return replay

    def optimize_pose(self, inputs, features, train_config=(True, True)):
        r_inputs, r_features = self._regroup(inputs, features)
        grasp_success = self([r_inputs, r_features])
        grasp_success = grasp_success.sum(dim=0, keepdim=True).transpose(0, 1)
        loss = -grasp_success
        grads = torch.autograd.grad(loss.sum(), [self.translations, self.rotations])
        for i, t_c in enumerate(train_config):
            if t_c:
                optimize(self.optimizer[i], self.pose_variables[i], [grads[i]], 1.0)
        self.loss_tracker.append(float(loss.detach().mean()))
        self.post_process()
        return {'loss': grasp_success.detach()}
