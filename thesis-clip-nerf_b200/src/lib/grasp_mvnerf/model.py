"""Drop-in for the reference's ``src/lib/grasp_mvnerf/model.py``: ``GraspMVNeRF`` -- the pose-query field with
the pose matrices as an input (``inputs = (transforms, images, K, E^-1)``) and a plain first-order train step
(reference :160-171)."""
import torch

from clipnerf_b200.trainer import allreduce_mean

from ..delta_ngf.model import _PoseQueryModel
from ..mvnerf.nerf_utils import optimize
from .layers import GraspReadout


class GraspMVNeRF(_PoseQueryModel):
    def __init__(self, n_features=256, original_image_size=(480, 640), n_points_train=2048, n_views=1, n_5d_poses=7,
                 softmax_before_loss=False, device=None, tf32_head=True, **kwargs):
        super().__init__()
        if n_features != 256:
            raise ValueError('the sm_100a path is specialised for n_features=256')
        self._init_common(n_views, n_5d_poses, GraspReadout(activation='elu'), tf32_head)
        self.n_points_train = n_points_train
        self.softmax_before_loss = softmax_before_loss
        self.original_image_size = tuple(original_image_size)
        self.loss_tracker = []
        if device is not None:
            self.to(device)

    def call(self, inputs, training=False, mask=None):
        feats = self.encode(self._to_dev(inputs[1]))
        return self._call(inputs, self.n_points_train, feats)

    def forward(self, inputs, training=False):
        return self.call(inputs, training=training)

    def infer(self, inputs, n_points_infer, batched_features):
        return self._call(inputs, n_points_infer, batched_features)

    def _call(self, inputs, n_points, batched_features):
        return self._query(inputs[1], inputs[2], inputs[3], inputs[0], n_points, batched_features)

    def train_step(self, data):
        inputs, labels = data
        readout_vars = list(self.grasp_readout.parameters())
        feats = self.encode(self._to_dev(inputs[1])).detach()
        energies = self._call(inputs, self.n_points_train, feats)
        if self.softmax_before_loss:
            energies = torch.softmax(energies, dim=-1)
        loss = self.loss(self._to_dev(labels), energies)
        gradients = torch.autograd.grad(loss, readout_vars, allow_unused=True)
        gradients = allreduce_mean(gradients, readout_vars, getattr(self, 'process_group', None))
        optimize(self.optimizer, readout_vars, gradients, 1.0)
        self.loss_tracker.append(float(loss.detach()))
        return {'loss': loss.detach()}
