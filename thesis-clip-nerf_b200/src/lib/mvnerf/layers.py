"""Drop-in for the hot-path layers of the reference's ``src/lib/mvnerf/layers.py`` (lines 262-411):
parameter containers with the Keras variable names / shapes / initialisers.  The arithmetic of
these layers runs inside the fused tcgen05 kernels (``cn_mlp_feature_fwd`` ...), which read one
flat fp32 vector per net; each module therefore owns a flat ``nn.Parameter`` and exposes the named
Keras variables as views into it."""
import math

import numpy as np
import torch

from clipnerf_b200 import params as P
from clipnerf_b200._abi import EMB_PARAMS, NET_PARAMS

_ACTS = ('relu', 'elu')


def _glorot_(t):
    lim = math.sqrt(6.0 / (t.shape[0] + t.shape[1]))
    with torch.no_grad():
        t.uniform_(-lim, lim)


class _FlatModule(torch.nn.Module):
    """``flat``: the module's variables as ONE fp32 vector.  Stand-alone it is the module's own ``nn.Parameter``;
    inside ``MVVNeRFRenderer`` it is a view into the model's ``[2, NET_PARAMS]`` parameter buffer (``share_storage``),
    so that the whole model trains through one flat buffer / one all-reduce."""
    _table = ()

    def __init__(self, numel):
        super().__init__()
        self.flat = torch.nn.Parameter(torch.zeros(numel))

    def share_storage(self, owner, attr, row, start):
        """give up the own parameter: ``flat`` becomes ``getattr(owner, attr)[row, start:start + numel]`` (resolved at
        every access, so ``owner.to(device)`` keeps working); the current values are NOT copied"""
        import weakref
        n = self._parameters['flat'].numel()
        del self._parameters['flat']
        self.__dict__['_view_of'] = (weakref.ref(owner), attr, row, start, n)

    def __getattr__(self, name):
        if name == 'flat' and '_view_of' in self.__dict__:
            owner, attr, row, start, n = self.__dict__['_view_of']
            return getattr(owner(), attr)[row, start:start + n]
        return super().__getattr__(name)

    def variables(self):
        """{keras variable name: view}"""
        return {n: self.flat[o:o + int(np.prod(s))].view(*s) for n, s, o in self._table}

    def reset_parameters(self):
        for name, v in self.variables().items():
            if name.endswith('kernel'):
                _glorot_(v)
            else:
                with torch.no_grad():
                    v.zero_()


class ResNetMLPBlock(torch.nn.Module):
    """y = shortcut(x) + act(act(x) W0 + b0) W1 + b1 (reference :262-298).  Stand-alone blocks
    (GraspReadout uses them with elu / a linear shortcut) run through torch; the six blocks of the
    NeRF embedding run inside the fused kernels."""

    def __init__(self, hidden_size, output_size, transform_shortcut=False, activation='relu', input_size=None):
        super().__init__()
        if activation not in _ACTS:
            raise ValueError(f'activation {activation} not supported')
        input_size = hidden_size if input_size is None else input_size
        self.layer_0 = torch.nn.Linear(input_size, hidden_size)
        self.layer_1 = torch.nn.Linear(hidden_size, output_size)
        self.shortcut = torch.nn.Linear(input_size, output_size, bias=False) if transform_shortcut else None
        self.act = torch.relu if activation == 'relu' else torch.nn.functional.elu

    def forward(self, inputs):
        r = self.layer_1(self.act(self.layer_0(self.act(inputs))))
        return (inputs if self.shortcut is None else self.shortcut(inputs)) + r


class MVResNetMLPNeRFEmbedding(_FlatModule):
    """Dense 379->128, 3 per-view blocks, mean over views, 3 fusion blocks (reference :334-379)."""
    _table = [(n.split('.', 1)[1], s, o) for n, s, o in P.variable_table() if n.startswith('embedding.')]

    def __init__(self, n_blocks=6, hidden_layer_size=128, n_views=2, n_freq=10, pos_encoding_freq=np.pi,
                 embed_direction_vector=False, complete_output=False):
        if n_blocks != 6 or hidden_layer_size != 128 or n_freq != 10 or not embed_direction_vector:
            raise ValueError('the sm_100a MLP kernels are specialised for MVResNetMLPNeRFEmbedding(6, 128, n_freq=10, '
                             'embed_direction_vector=True), the only configuration the reference models build')
        super().__init__(EMB_PARAMS)
        self.n_views, self.n_freq, self.pos_encoding_freq = n_views, n_freq, pos_encoding_freq
        self.embed_direction_vector, self.complete_output = embed_direction_vector, complete_output
        self.reset_parameters()

    def forward(self, inputs):
        raise NotImplementedError('stand-alone embedding calls are not part of the render path: '
                                  'MVVNeRFRenderer._call runs the embedding inside the fused render kernels')


class RenderReadout(_FlatModule):
    """relu -> Dense 128->4 -> (sigmoid rgb, softplus sigma) (reference :382-397)."""
    _table = [(n.split('.', 1)[1], s, o - EMB_PARAMS) for n, s, o in P.variable_table() if n.startswith('readout.')]

    def __init__(self, output_size=4):
        if output_size != 4:
            raise ValueError('RenderReadout is specialised for 4 outputs (rgb + sigma)')
        super().__init__(NET_PARAMS - EMB_PARAMS)
        self.reset_parameters()

    def forward(self, inputs):
        v = self.variables()
        y = torch.relu(inputs) @ v['output_layer.kernel'] + v['output_layer.bias']
        return torch.sigmoid(y[..., :3]), torch.nn.functional.softplus(y[..., 3])


class Readout(torch.nn.Module):
    """relu -> Dense (reference :400-411); used by the affordance head."""

    def __init__(self, output_size, use_bias=True, input_size=64):
        super().__init__()
        self.output_layer = torch.nn.Linear(input_size, output_size, bias=use_bias)

    def forward(self, inputs):
        return self.output_layer(torch.relu(inputs))
