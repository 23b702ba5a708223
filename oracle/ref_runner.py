"""Runs the reference's OWN source (``/root/reference/src/lib/mvnerf/*.py``, unmodified) on the
torch-backed TensorFlow stand-in of ``oracle/tf_shim``.  TEST INFRASTRUCTURE ONLY; works only where
the reference checkout exists (the build container)."""
from __future__ import annotations

import types
from typing import Dict

import torch

from . import tf_shim


def _set_dense(dense, kernel, bias):
    dense.kernel = kernel
    dense.bias = bias


def build_reference_nets(ly, p: Dict[str, torch.Tensor], n_views: int):
    """reference ``MVResNetMLPNeRFEmbedding`` / ``RenderReadout`` objects carrying the oracle's parameters"""
    nets = {}
    for net in ('coarse', 'fine'):
        emb = ly.MVResNetMLPNeRFEmbedding(6, 128, embed_direction_vector=True, n_views=n_views)
        pre = f'{net}_embedding'
        _set_dense(emb.layer_0, p[f'{pre}.layer_0.kernel'], p[f'{pre}.layer_0.bias'])
        for grp, blocks in (('feature_blocks', emb.feature_blocks), ('fusion_blocks', emb.fusion_blocks)):
            for i, blk in enumerate(blocks):
                for l in ('layer_0', 'layer_1'):
                    _set_dense(getattr(blk, l), p[f'{pre}.{grp}.{i}.{l}.kernel'], p[f'{pre}.{grp}.{i}.{l}.bias'])
        ro = ly.RenderReadout(4)
        _set_dense(ro.output_layer, p[f'{net}_readout.output_layer.kernel'], p[f'{net}_readout.output_layer.bias'])
        nets[net] = (emb, ro)
    return nets


def reference_call(inputs, combined_features, p, *, near, far, n_samples, u_strat, u_pdf):
    """``MVVNeRFRenderer._call`` of the reference (model_v0.py:113-184), executed from its source, with the
    two ``tf.random.uniform`` draws (nerf_utils.py:57,151) supplied explicitly."""
    nu, ly, mv = tf_shim.load_reference()
    n_views = inputs[2].shape[1]
    nets = build_reference_nets(ly, p, n_views)
    fake_self = types.SimpleNamespace(near=near, far=far, n_samples=n_samples, n_views=n_views,
                                      coarse_embedding=nets['coarse'][0], coarse_readout=nets['coarse'][1],
                                      fine_embedding=nets['fine'][0], fine_readout=nets['fine'][1])
    b, nr = inputs[0].shape[:2]
    tf_shim.queue_uniform(u_strat, u_pdf)
    out = mv.MVVNeRFRenderer._call(fake_self, tuple(inputs), nr, b, combined_features)
    assert not tf_shim._UNIFORM_QUEUE, 'the reference did not consume both random draws'
    return out
