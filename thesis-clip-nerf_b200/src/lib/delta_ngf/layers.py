"""Drop-in for the reference's ``src/lib/delta_ngf/layers.py`` (= ``lmvnerf/layers.py``) and, with
``activation=``, ``grasp_mvnerf/layers.py``: the affordance head ``GraspReadout``.

4 x [b, np, 42, 128] stage outputs -> Dense 128->64 (+act) each -> concat 256 -> Dense 64 (+act) ->
flatten (42, 64) = 2688 -> ResNetMLPBlock(2688->128->64, linear shortcut) -> ResNetMLPBlock(64->64->64) ->
relu -> Dense 64->1 (reference delta_ngf/layers.py:8-42).  582 337 parameters.

The head is ~98 kFLOP per query point plus ~1 MFLOP per pose of plain dense layers on tensors the fused
kernels hand over in fp32; it runs as torch (cuBLAS) modules so that torch autograd provides the double
backward ``DeltaNGF.train_step`` takes through it (delta_ngf/model.py:264-296)."""
import math

import torch

from ..mvnerf.layers import Readout, ResNetMLPBlock

_ACT = {'elu': torch.nn.functional.elu, 'relu': torch.relu}


class _tf32_matmul:
    """scoped TensorFloat-32 for torch's fp32 matmuls (the process-wide switch is restored on exit)"""

    def __enter__(self):
        self.old = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = True

    def __exit__(self, *a):
        torch.backends.cuda.matmul.allow_tf32 = self.old


class _LinearTF32(torch.autograd.Function):
    """y = x W^T + b with the three matmuls (forward, d/dx, d/dW) in TensorFloat-32.  TensorFlow, which the
    reference runs on, executes fp32 matmuls in TF32 on Ampere-and-later GPUs by default
    (tf.config.experimental.enable_tensor_float_32_execution), so this is the arithmetic of the reference's own
    GraspReadout on such a GPU; it takes the head from 4 ms to well under 1 ms per optimize_pose step at the
    validation size.  The backward is made of differentiable torch ops, so the double backward of
    DeltaNGF.train_step goes through."""

    @staticmethod
    def forward(ctx, x, w, b):
        ctx.save_for_backward(x, w)
        ctx.has_bias = b is not None
        with _tf32_matmul():
            return torch.nn.functional.linear(x, w, b)

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        gx = gw = gb = None
        g2 = gy.reshape(-1, gy.shape[-1])
        with _tf32_matmul():
            if ctx.needs_input_grad[0]:
                gx = gy @ w
            if ctx.needs_input_grad[1]:
                gw = g2.t() @ x.reshape(-1, x.shape[-1])
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = g2.sum(0)
        return gx, gw, gb


def _linear(layer, x, tf32):
    if tf32 and x.is_cuda:
        return _LinearTF32.apply(x, layer.weight, layer.bias)
    return layer(x)


def _he_normal_(linear):
    with torch.no_grad():
        linear.weight.normal_(0.0, math.sqrt(2.0 / linear.in_features))
        if linear.bias is not None:
            linear.bias.zero_()


class GraspReadout(torch.nn.Module):
    def __init__(self, use_bias=False, activation='elu', **kwargs):
        super().__init__()
        if activation not in _ACT:
            raise ValueError(f'activation {activation} not supported')
        self.act = _ACT[activation]
        self.activation = activation
        self._pack_cache = None           # (versions, packed operand images) of the fused inference kernels
        self.tf32 = False                 # set by the models (``tf32_head``); plain fp32 when used stand-alone
        for i in range(1, 5):
            setattr(self, f'activation_downscale_{i}', torch.nn.Linear(128, 64))
        self.combined_activation_downscale = torch.nn.Linear(256, 64)
        self.readout = torch.nn.Sequential(
            ResNetMLPBlock(128, 64, transform_shortcut=True, activation=activation, input_size=42 * 64),
            ResNetMLPBlock(64, 64, activation=activation),
            Readout(1, use_bias=use_bias, input_size=64))
        self.reset_parameters()

    def reset_parameters(self):
        """he_normal kernels, zero biases (delta_ngf/layers.py:12-28); the combined downscale keeps the
        Keras default (glorot_uniform)."""
        for m in self.modules():
            if isinstance(m, torch.nn.Linear) and m is not self.combined_activation_downscale:
                _he_normal_(m)
        torch.nn.init.xavier_uniform_(self.combined_activation_downscale.weight)
        torch.nn.init.zeros_(self.combined_activation_downscale.bias)

    def forward(self, inputs, *args, **kwargs):
        t = self.tf32
        ds = [self.act(_linear(getattr(self, f'activation_downscale_{i + 1}'), inputs[i], t)) for i in range(4)]
        c = self.act(_linear(self.combined_activation_downscale, torch.cat(ds, dim=-1), t))
        c = c.reshape(*c.shape[:-2], -1)                      # 'b np n5 d -> b np (n5 d)'
        for blk in (self.readout[0], self.readout[1]):        # ResNetMLPBlock (mvnerf/layers.py:262-298)
            r = _linear(blk.layer_1, blk.act(_linear(blk.layer_0, blk.act(c), t)), t)
            c = (c if blk.shortcut is None else _linear(blk.shortcut, c, t)) + r
        return _linear(self.readout[2].output_layer, torch.relu(c), t)[..., 0]

    # ---- fused inference path (cn_grasp_head_fwd / _bwd): frozen weights, gradient to the inputs only
    def flat_keras(self):
        """the variables as ONE fp32 vector in Keras order, kernel[in][out] (include/clipnerf_b200.h CN_HEAD_PARAMS)"""
        named = self.named_keras_variables()
        parts = [v.detach().reshape(-1) for v in named.values()]
        if 'grasp_readout.readout.2.output_layer.bias' not in named:
            parts.append(parts[0].new_zeros(1))
        return torch.cat(parts).to(torch.float32)

    def packed(self):
        """bf16 operand images of the head's matrices, rebuilt only when a variable changed (tensor version counters)"""
        from clipnerf_b200 import ops
        params = list(self.parameters())
        key = tuple((p.data_ptr(), p._version) for p in params)
        if self._pack_cache is None or self._pack_cache[0] != key:
            self._pack_cache = (key, ops.grasp_head_pack(self.flat_keras()))
        return self._pack_cache[1]

    def fused_available(self, n_offsets, grad_to_weights):
        """the tcgen05 head covers 42 query offsets per pose and returns no parameter gradients"""
        w0 = self.readout[0].layer_0.weight
        return w0.is_cuda and n_offsets == 42 and w0.shape[1] == 42 * 64 and not grad_to_weights

    # ---- Keras-named view of the variables (kernel[in, out]), for parity tests and checkpoint interchange
    def named_keras_variables(self, prefix='grasp_readout', grad=False):
        """{keras name: tensor}; ``grad=True`` returns the ``.grad`` of every variable in the same (kernel[in, out]) form"""
        def kern(l):
            return (l.weight.grad if grad else l.weight).t()

        def bias(l):
            return l.bias.grad if grad else l.bias
        out = {}
        for i in range(1, 5):
            l = getattr(self, f'activation_downscale_{i}')
            out[f'{prefix}.activation_downscale_{i}.kernel'], out[f'{prefix}.activation_downscale_{i}.bias'] = kern(l), bias(l)
        l = self.combined_activation_downscale
        out[f'{prefix}.combined_activation_downscale.kernel'], out[f'{prefix}.combined_activation_downscale.bias'] = kern(l), bias(l)
        for j in (0, 1):
            blk = self.readout[j]
            for ln in ('layer_0', 'layer_1'):
                l = getattr(blk, ln)
                out[f'{prefix}.readout.{j}.{ln}.kernel'], out[f'{prefix}.readout.{j}.{ln}.bias'] = kern(l), bias(l)
            if blk.shortcut is not None:
                out[f'{prefix}.readout.{j}.shortcut.kernel'] = kern(blk.shortcut)
        l = self.readout[2].output_layer
        out[f'{prefix}.readout.2.output_layer.kernel'] = kern(l)
        if l.bias is not None:
            out[f'{prefix}.readout.2.output_layer.bias'] = bias(l)
        return out

    def load_keras_variables(self, named, prefix='grasp_readout'):
        with torch.no_grad():
            for name, v in self.named_keras_variables(prefix).items():
                v.copy_(named[name].to(v.device))
