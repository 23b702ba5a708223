"""Drop-in for ``src/lib/grasp_mvnerf/layers.py``: ``GraspReadout(use_bias=True, activation='relu')`` --
the same head with a selectable activation and Keras-default (glorot_uniform) kernels."""
import torch

from ..delta_ngf.layers import GraspReadout as _GraspReadout


class GraspReadout(_GraspReadout):
    def __init__(self, use_bias=True, activation='relu', **kwargs):
        super().__init__(use_bias=use_bias, activation=activation)

    def reset_parameters(self):
        for m in self.modules():
            if isinstance(m, torch.nn.Linear):
                torch.nn.init.xavier_uniform_(m.weight)
                if m.bias is not None:
                    torch.nn.init.zeros_(m.bias)
