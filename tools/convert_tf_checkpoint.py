#!/usr/bin/env python
"""Reference checkpoint -> ``.npz`` in the variable naming of ``src/lib/mvnerf/encoders.py`` (and of the flat NeRF nets).

Runs in the REFERENCE's environment (TensorFlow 2.11 + the reference repository on PYTHONPATH), not here: TensorFlow is
not installed in the build container, so this script is exercised only by its name resolver's unit logic (the mapping
of attribute paths below); the reader side (``encoders.load_keras_variables``) is tested in ``tests/test_encoders.py``.

    python tools/convert_tf_checkpoint.py <reference checkpoint prefix> <out.npz> [--model v4]

The reference stores one Keras ``save_weights`` checkpoint per sub-module (``model_v0.py:199-211``:
``<prefix>_{coarse_embedding,coarse_readout,fine_embedding,fine_readout,visual_features,combine_clip_visual}``).
The script builds the reference model, lets ``model.load(prefix)`` restore it, and then walks the SAME attribute
paths the torch mirrors use -- variable identity comes from the object structure, not from TensorFlow's checkpoint
key strings.  Layouts are Keras' (HWIO conv kernels, ``[in, out]`` Dense kernels): the loader permutes."""
import argparse
import re
import sys

import numpy as np

_MHA = {'query': '_query_dense', 'key': '_key_dense', 'value': '_value_dense', 'output': '_output_dense'}


def resolve(root, dotted):
    """the Keras variable behind one of ``encoders.keras_variables``' names, e.g.
    'vision_transformer.vit.transformer_blocks.2.0.mlp.2.kernel' -> root.vision_transformer.vit.transformer_blocks[2]
    .layers[0].mlp.layers[1].kernel (torch's Sequential counts its GELU module, Keras' does not)"""
    obj, parts = root, dotted.split('.')
    for i, t in enumerate(parts):
        prev = parts[i - 1] if i else ''
        if t.isdigit():
            k = int(t)
            if prev == 'mlp':
                k = {0: 0, 2: 1}[k]
            obj = obj.layers[k] if hasattr(obj, 'layers') else obj[k]
        elif t in _MHA and prev == 'attention':
            obj = getattr(obj, _MHA[t])
        elif t == 'dense' and re.match(r'multiply_fusion_\d', prev):
            obj = obj.tile.dense
        else:
            obj = getattr(obj, t)
    return obj


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('prefix')
    ap.add_argument('out')
    ap.add_argument('--model', default='v4', choices=['v0', 'v1', 'v2', 'v3', 'v4'])
    ap.add_argument('--names', required=True, help='text file with one encoders.keras_variables name per line '
                                                   '(python -c "from src.lib.mvnerf import encoders as E; ..." in this repo)')
    args = ap.parse_args()
    import importlib
    ref = importlib.import_module(f'src.lib.mvnerf.model_{args.model}')          # the REFERENCE package
    model = ref.MVVNeRFRenderer(512, 512, n_views=3)
    if not model.load(args.prefix):
        sys.exit(f'checkpoint {args.prefix}_* incomplete')
    out = {}
    for line in open(args.names):
        name = line.strip()
        if not name:
            continue
        group, rest = name.split('.', 1)                                         # visual_features.* / combine_clip_visual.*
        out[name] = np.asarray(resolve(getattr(model, group), rest).numpy())
    for net in ('coarse', 'fine'):                                               # the NeRF nets: Keras order, kernel[in][out]
        for sub in ('embedding', 'readout'):
            for i, v in enumerate(getattr(model, f'{net}_{sub}').variables):       # Keras creation order = flat layout order
                out[f'{net}_{sub}/{i:03d}'] = v.numpy()
    np.savez(args.out, **out)
    print(f'wrote {len(out)} arrays to {args.out}')


if __name__ == '__main__':
    main()
