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


def _set_block(blk, p, pre):
    for l in ('layer_0', 'layer_1'):
        _set_dense(getattr(blk, l), p[f'{pre}.{l}.kernel'], p[f'{pre}.{l}.bias'])
    if blk.shortcut is not None:
        _set_dense(blk.shortcut, p[f'{pre}.shortcut.kernel'], None)


def build_reference_query_model(p: Dict[str, torch.Tensor], n_views: int, batch_size: int, n_points: int,
                                which: str = 'delta_ngf', rotation_representation: str = 'quaternion'):
    """the reference's ``DeltaNGF`` / ``GraspMVNeRF`` object (constructed by its own ``__init__``, so
    ``transforms_to_check`` is the reference's) carrying the oracle's parameters.  The image encoder
    (``VisualFeatures``: ViT + conv, outside the path -- its output is an input here) is stubbed out."""
    md, _ly = tf_shim.load_reference_query(which)
    md.VisualFeatures = lambda *a, **k: None
    if which == 'delta_ngf':
        model = md.DeltaNGF(n_views=n_views, batch_size=batch_size, n_points_train=n_points,
                            rotation_representation=rotation_representation)
    else:
        model = md.GraspMVNeRF(n_views=n_views, n_points_train=n_points)
    emb, pre = model.fine_embedding, 'fine_embedding'
    _set_dense(emb.layer_0, p[f'{pre}.layer_0.kernel'], p[f'{pre}.layer_0.bias'])
    for grp, blocks in (('feature_blocks', emb.feature_blocks), ('fusion_blocks', emb.fusion_blocks)):
        for i, blk in enumerate(blocks):
            _set_block(blk, p, f'{pre}.{grp}.{i}')
    gr, pre = model.grasp_readout, 'grasp_readout'
    for i in range(1, 5):
        _set_dense(getattr(gr, f'activation_downscale_{i}'), p[f'{pre}.activation_downscale_{i}.kernel'],
                   p[f'{pre}.activation_downscale_{i}.bias'])
    _set_dense(gr.combined_activation_downscale, p[f'{pre}.combined_activation_downscale.kernel'],
               p[f'{pre}.combined_activation_downscale.bias'])
    seq = gr.readout.layers
    _set_block(seq[0], p, f'{pre}.readout.0')
    _set_block(seq[1], p, f'{pre}.readout.1')
    _set_dense(seq[2].output_layer, p[f'{pre}.readout.2.output_layer.kernel'],
               p.get(f'{pre}.readout.2.output_layer.bias'))
    seq[2].output_layer.use_bias = f'{pre}.readout.2.output_layer.bias' in p
    return model


def reference_query_call(images, K, Einv, transforms, features, p, which: str = 'delta_ngf'):
    """``DeltaNGF._call`` (delta_ngf/model.py:188-245) / ``GraspMVNeRF._call`` (grasp_mvnerf/model.py:113-158)
    executed from the reference's source."""
    b, nv = images.shape[:2]
    n_points = transforms.shape[1]
    model = build_reference_query_model(p, nv, b, n_points, which)
    if which == 'delta_ngf':
        return model._call((None, None, None, None, images, K, Einv), transforms, n_points, features)
    return model._call((transforms, images, K, Einv), n_points, features)


def reference_compute_matrices(translations, rotations, rotation_representation):
    """``DeltaNGF.compute_matrices`` (delta_ngf/model.py:172-183) from the reference's source"""
    md, _ = tf_shim.load_reference_query('delta_ngf')
    fake = types.SimpleNamespace(rotation_representation=rotation_representation, translations=translations,
                                 quaternions=rotations, six_d_rotations=rotations)
    return md.DeltaNGF.compute_matrices(fake)
