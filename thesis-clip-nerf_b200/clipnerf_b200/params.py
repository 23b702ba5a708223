"""Flat fp32 parameter layout of one NeRF net (MVResNetMLPNeRFEmbedding(6,128) + RenderReadout(4),
reference ``src/lib/mvnerf/layers.py:334-397``) and its mapping to Keras variable names.

The CUDA path reads ONE contiguous vector per net (``cn_pack_weights``) and writes gradients into a
vector of the same layout, so a single NCCL all-reduce covers every trainable NeRF variable."""
from __future__ import annotations

import math
from typing import Dict, List, Tuple

import numpy as np
import torch

from ._abi import EMB_PARAMS, NET_PARAMS

IN_DIM = 379
HIDDEN = 128


def variable_table() -> List[Tuple[str, Tuple[int, ...], int]]:
    """[(name, shape, float offset)] in flat order.  Names follow the Keras attributes of the
    reference classes; ``embedding.*`` belongs to ``coarse_embedding`` / ``fine_embedding`` and
    ``readout.*`` to ``coarse_readout`` / ``fine_readout`` (checkpoint groups of model_v0.py:199-211)."""
    out = []
    off = 0

    def add(name, shape):
        nonlocal off
        out.append((name, shape, off))
        off += int(np.prod(shape))

    add('embedding.layer_0.kernel', (IN_DIM, HIDDEN))
    add('embedding.layer_0.bias', (HIDDEN,))
    for grp in ('feature_blocks', 'fusion_blocks'):
        for i in range(3):
            for l in ('layer_0', 'layer_1'):
                add(f'embedding.{grp}.{i}.{l}.kernel', (HIDDEN, HIDDEN))
                add(f'embedding.{grp}.{i}.{l}.bias', (HIDDEN,))
    assert off == EMB_PARAMS
    add('readout.output_layer.kernel', (HIDDEN, 4))
    add('readout.output_layer.bias', (4,))
    assert off == NET_PARAMS
    return out


_TABLE = variable_table()


def views(flat: torch.Tensor) -> Dict[str, torch.Tensor]:
    """Named views into a flat parameter (or gradient) vector."""
    assert flat.numel() == NET_PARAMS
    return {name: flat[off:off + int(np.prod(shape))].view(*shape) for name, shape, off in _TABLE}


def glorot_init(generator: torch.Generator = None, bias_std: float = 0.0) -> torch.Tensor:
    """Keras defaults: glorot_uniform kernels, zero biases (``bias_std`` > 0 for tests)."""
    flat = torch.zeros(NET_PARAMS, dtype=torch.float32)
    for name, shape, off in _TABLE:
        n = int(np.prod(shape))
        if name.endswith('kernel'):
            lim = math.sqrt(6.0 / (shape[0] + shape[1]))
            flat[off:off + n] = ((torch.rand(n, generator=generator) * 2 - 1) * lim)
        elif bias_std > 0:
            flat[off:off + n] = torch.randn(n, generator=generator) * bias_std
    return flat


def from_named(named: Dict[str, torch.Tensor], net: str) -> torch.Tensor:
    """Build the flat vector of net ``'coarse'`` / ``'fine'`` from a dict keyed like the oracle's
    ``{net}_embedding.<var>`` / ``{net}_readout.<var>``."""
    flat = torch.empty(NET_PARAMS, dtype=torch.float32)
    for name, shape, off in _TABLE:
        grp, rest = name.split('.', 1)
        src = named[f'{net}_{grp}.{rest}']
        flat[off:off + src.numel()] = src.reshape(-1).to(torch.float32)
    return flat


def to_named(flat: torch.Tensor, net: str) -> Dict[str, torch.Tensor]:
    return {f"{net}_{name.split('.', 1)[0]}.{name.split('.', 1)[1]}": v for name, v in views(flat).items()}
