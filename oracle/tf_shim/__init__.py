"""A small torch-backed stand-in for the parts of TensorFlow 2.x / Keras / tensorflow_addons that the
reference's hot-path source files touch.  TEST INFRASTRUCTURE ONLY (see oracle/nerf_oracle.py header).

Purpose: TensorFlow cannot be installed here, so the reference cannot run as shipped.  With this
stand-in registered as ``tensorflow`` the reference's OWN files
(``src/lib/mvnerf/nerf_utils.py``, ``layers.py``, ``model_v0.py`` -- unmodified, read from
``/root/reference``) execute on torch CPU tensors, which pins the oracle restatement against the
reference's code: tensor layouts, rearranges, formulas and quirks.  What it cannot pin is the
arithmetic *inside* TF kernels; each op below is the documented TF semantics:

* ``tf.gather_nd`` out-of-range index -> 0 (GPU kernel behaviour; the CPU kernel raises);
* ``tf.matmul`` -> ``torch.matmul`` (batch broadcasting identical; summation order unspecified in both);
* ``tf.math.cumsum / cumprod(exclusive=...)``, ``tf.scan``, ``tf.sort`` -> their definitions;
* ``tf.random.uniform`` pops tensors queued with ``queue_uniform`` so draws are explicit;
* ``tensorflow_addons.image.interpolate_bilinear`` is third-party and NOT part of the reference tree:
  it maps to the oracle's restatement of the TFA algorithm (so it is not independently pinned).

``install()`` registers the modules; ``load_reference()`` imports the reference package under the
alias ``refsrc`` (its ``src`` directory) so it cannot collide with this repo's own ``src`` mirror."""
from __future__ import annotations

import importlib
import os
import sys
import types

import numpy as np
import torch

REFERENCE_ROOT = os.environ.get('CLIPNERF_REFERENCE_ROOT', '/root/reference')

_UNIFORM_QUEUE = []


def queue_uniform(*tensors):
    """tensors returned (in order) by the next calls of tf.random.uniform"""
    _UNIFORM_QUEUE.extend(tensors)


def _dtype(d):
    if d is None or isinstance(d, torch.dtype):
        return d
    return {'float32': torch.float32, 'float64': torch.float64, 'int32': torch.int32, 'int64': torch.int64,
            'bool': torch.bool}[str(d)]


def _t(x, like=None):
    if torch.is_tensor(x):
        return x
    dt = like.dtype if (like is not None and like.is_floating_point() and isinstance(x, float)) else None
    return torch.as_tensor(x, dtype=dt)


def _ints(shape):
    return [int(s) for s in shape]


def _build_tf():
    tf = types.ModuleType('tensorflow')
    tf.float32, tf.float64, tf.int32, tf.int64, tf.bool = torch.float32, torch.float64, torch.int32, torch.int64, torch.bool
    tf.newaxis = None
    # dummy classes: einops probes `tensorflow.Tensor / Variable`; torch tensors must stay with its torch backend
    tf.Tensor = type('Tensor', (), {})

    class Variable:
        """tf.Variable(value, trainable=...) -> a leaf torch tensor (isinstance checks against this class stay False)"""
        def __new__(cls, value, trainable=True, dtype=None, **k):
            t = torch.as_tensor(np.asarray(value), dtype=_dtype(dtype)).clone()
            return t.requires_grad_(bool(trainable) and t.is_floating_point())
    tf.Variable = Variable
    tf.tile = lambda x, multiples: x.repeat(*_ints(multiples))
    tf.reshape = lambda x, shape: x.reshape(_ints(shape))
    tf.linalg = types.SimpleNamespace(
        normalize=lambda x, axis=-1: (x / x.norm(dim=axis, keepdim=True), x.norm(dim=axis, keepdim=True)),
        cross=lambda a, b: torch.linalg.cross(a, b, dim=-1))

    def convert_to_tensor(x, dtype=None):
        return torch.as_tensor(np.asarray(x), dtype=_dtype(dtype)) if not torch.is_tensor(x) else x.to(_dtype(dtype) or x.dtype)
    tf.convert_to_tensor = convert_to_tensor
    tf.constant = lambda x, dtype=None: convert_to_tensor(x, dtype)
    tf.ones_like = torch.ones_like
    tf.zeros_like = lambda x, dtype=None: torch.zeros_like(x, dtype=_dtype(dtype))
    tf.ones = lambda shape, dtype=torch.float32: torch.ones(_ints(shape), dtype=_dtype(dtype))
    tf.zeros = lambda shape, dtype=torch.float32: torch.zeros(_ints(shape), dtype=_dtype(dtype))
    tf.shape = lambda x: list(x.shape)
    tf.concat = lambda xs, axis: torch.cat([_t(x) for x in xs], dim=axis)
    tf.stack = lambda xs, axis=0: torch.stack(list(xs), dim=axis)
    tf.matmul = lambda a, b: torch.matmul(a.to(torch.promote_types(a.dtype, b.dtype)), b.to(torch.promote_types(a.dtype, b.dtype)))
    tf.maximum = lambda a, b: torch.maximum(_t(a, b if torch.is_tensor(b) else None), _t(b, a if torch.is_tensor(a) else None))
    tf.minimum = lambda a, b: torch.minimum(_t(a, b if torch.is_tensor(b) else None), _t(b, a if torch.is_tensor(a) else None))
    tf.clip_by_value = lambda x, clip_value_min, clip_value_max: torch.clamp(x, min=clip_value_min, max=clip_value_max)
    tf.pow = lambda a, b: torch.pow(_t(a, b if torch.is_tensor(b) else None), b)
    tf.range = lambda n, dtype='int32': torch.arange(n, dtype=_dtype(dtype))
    tf.sin, tf.cos, tf.exp = torch.sin, torch.cos, torch.exp
    tf.reduce_sum = lambda x, axis=None, keepdims=False: torch.sum(x, dim=axis, keepdim=keepdims) if axis is not None else x.sum()
    tf.reduce_mean = lambda x, axis=None, keepdims=False: torch.mean(x, dim=axis, keepdim=keepdims) if axis is not None else x.mean()
    tf.where = lambda c, a, b: torch.where(c, a, b)
    tf.transpose = lambda x, perm: x.permute(*perm)
    tf.cast = lambda x, dtype: _t(x).to(_dtype(dtype))
    tf.greater_equal = lambda a, b: a >= b
    tf.expand_dims = lambda x, axis: x.unsqueeze(axis)
    tf.sort = lambda x, axis=-1: torch.sort(x, dim=axis, stable=True).values
    tf.cond = lambda pred, a, b: a() if bool(pred) else b()

    def scan(fn, elems, initializer):
        acc, outs = initializer, []
        for i in range(elems.shape[0]):
            acc = fn(acc, elems[i])
            outs.append(acc)
        return torch.stack(outs, dim=0)
    tf.scan = scan

    def gather_nd(params, indices, batch_dims=0):
        assert batch_dims == params.dim() - 1 and indices.shape[-1] == 1, 'only the form used by sample_pdf'
        idx = indices[..., 0].to(torch.int64)
        n = params.shape[-1]
        ok = (idx >= 0) & (idx < n)
        val = torch.gather(params, -1, idx.clamp(0, n - 1)[..., None])[..., 0]
        return torch.where(ok, val, torch.zeros_like(val))          # out of range -> 0 (GPU kernel)
    tf.gather_nd = gather_nd

    tf.math = types.SimpleNamespace(
        abs=torch.abs, divide=lambda a, b: a / b,
        maximum=lambda a, b: torch.maximum(_t(a, b if torch.is_tensor(b) else None), _t(b, a if torch.is_tensor(a) else None)),
        cumsum=lambda x, axis=0, exclusive=False: _cum(torch.cumsum, x, axis, exclusive, 0.0),
        cumprod=lambda x, axis=0, exclusive=False: _cum(torch.cumprod, x, axis, exclusive, 1.0))

    def uniform(shape, minval=0, maxval=1, dtype=torch.float32):
        if _UNIFORM_QUEUE:
            u = _UNIFORM_QUEUE.pop(0)
            assert list(u.shape) == _ints(shape), (u.shape, shape)
            return u
        return torch.rand(_ints(shape)) * (maxval - minval) + minval
    tf.random = types.SimpleNamespace(uniform=uniform, normal=lambda shape, **k: torch.randn(_ints(shape)))

    def function(*args, **kwargs):
        if len(args) == 1 and callable(args[0]) and not kwargs:
            return args[0]
        return lambda f: f
    tf.function = function
    tf.TensorSpec = lambda *a, **k: None
    tf.config = types.SimpleNamespace(list_physical_devices=lambda *a: [], experimental=types.SimpleNamespace(
        set_memory_growth=lambda *a: None))

    # ---- keras
    class Layer:
        def __init__(self, *a, **k):
            pass

        def __call__(self, *a, **k):
            return self.call(*a, **k)

    class Dense(Layer):
        def __init__(self, units, use_bias=True, kernel_initializer='glorot_uniform', activation=None, **k):
            self.units, self.use_bias, self.kernel, self.bias = units, use_bias, None, None
            self.activation = {None: None, 'elu': torch.nn.functional.elu, 'relu': torch.relu}[activation]

        def build(self, in_dim, dtype=torch.float32):
            lim = (6.0 / (in_dim + self.units)) ** 0.5
            self.kernel = (torch.rand(in_dim, self.units, dtype=dtype) * 2 - 1) * lim
            self.bias = torch.zeros(self.units, dtype=dtype) if self.use_bias else None

        def call(self, x):
            if self.kernel is None:
                self.build(x.shape[-1], x.dtype)
            y = x @ self.kernel
            y = y + self.bias if self.use_bias else y
            return y if self.activation is None else self.activation(y)

    class _Act(Layer):
        def __init__(self, fn):
            self.fn = fn

        def call(self, x):
            return self.fn(x)

    def _unsupported(name):
        class U(Layer):
            """constructible (model constructors build their encoders), not callable"""
            def call(self, *a, **k):
                raise NotImplementedError(f'tf_shim: keras.layers.{name} is outside the hot path')
        U.__name__ = name
        return U

    class Sequential(Layer):
        def __init__(self, layers=None, **k):
            self.layers = list(layers or [])

        def call(self, x, *a, **k):
            for l in self.layers:
                x = l(x)
            return x

    acts = types.SimpleNamespace(relu=torch.relu, softplus=torch.nn.functional.softplus, sigmoid=torch.sigmoid,
                                 elu=torch.nn.functional.elu)
    layers = types.SimpleNamespace(Layer=Layer, Dense=Dense, ReLU=lambda: _Act(torch.relu),
                                   ELU=lambda: _Act(torch.nn.functional.elu), Activation=lambda fn: _Act(fn))
    for n in ('Conv2D', 'BatchNormalization', 'MultiHeadAttention', 'LayerNormalization', 'UpSampling2D', 'Resizing',
              'MaxPool2D', 'Dropout', 'Add', 'Concatenate', 'Conv2DTranspose'):
        setattr(layers, n, _unsupported(n))
    keras = types.SimpleNamespace(
        Model=Layer, layers=layers, activations=acts, Sequential=Sequential,
        optimizers=types.SimpleNamespace(schedules=types.SimpleNamespace(LearningRateSchedule=object), Adam=object),
        utils=types.SimpleNamespace(Sequence=object),
        metrics=types.SimpleNamespace(Mean=lambda name=None: None),
        losses=types.SimpleNamespace(MeanSquaredError=lambda: (lambda y, p: torch.mean((p - y) ** 2)),
                                     CosineSimilarity=lambda axis=-1: None),
        models=types.SimpleNamespace(load_model=lambda *a, **k: (_ for _ in ()).throw(NotImplementedError('CLIP weights'))))
    tf.keras = keras
    return tf


def _cum(fn, x, axis, exclusive, unit):
    incl = fn(x, dim=axis)
    if not exclusive:
        return incl
    first = torch.full_like(incl.narrow(axis, 0, 1), unit)
    return torch.cat([first, incl.narrow(axis, 0, x.shape[axis] - 1)], dim=axis)


def install():
    """register the stand-ins (idempotent); refuses to shadow a real TensorFlow"""
    if 'tensorflow' in sys.modules and not getattr(sys.modules['tensorflow'], '_is_clipnerf_shim', False):
        raise RuntimeError('a real tensorflow is importable; run the reference directly instead of the shim')
    tf = _build_tf()
    tf._is_clipnerf_shim = True
    sys.modules['tensorflow'] = tf
    for name in ('tensorflow.python', 'tensorflow.python.ops', 'tensorflow.python.ops.math_ops'):
        sys.modules[name] = types.ModuleType(name)
    from oracle import nerf_oracle as O
    tfa = types.ModuleType('tensorflow_addons')
    tfa.image = types.ModuleType('tensorflow_addons.image')
    tfa.image.interpolate_bilinear = lambda grid, query_points, indexing='ij': O.interpolate_bilinear(
        grid, query_points, indexing=indexing)
    tfa.optimizers = types.ModuleType('tensorflow_addons.optimizers')
    tfa.optimizers.MultiOptimizer = object
    sys.modules['tensorflow_addons'] = tfa
    sys.modules['tensorflow_addons.image'] = tfa.image
    sys.modules['tensorflow_addons.optimizers'] = tfa.optimizers
    # tensorflow_graphics (third party, not in the reference tree): only rotation_matrix_3d.from_quaternion is
    # used (delta_ngf/model.py:29); it maps to the oracle's restatement of the published formula
    tfg = types.ModuleType('tensorflow_graphics')
    tfg.geometry = types.ModuleType('tensorflow_graphics.geometry')
    tfg.geometry.transformation = types.ModuleType('tensorflow_graphics.geometry.transformation')
    tfg.geometry.transformation.rotation_matrix_3d = types.SimpleNamespace(from_quaternion=O.rotation_from_quaternion)
    sys.modules['tensorflow_graphics'] = tfg
    sys.modules['tensorflow_graphics.geometry'] = tfg.geometry
    sys.modules['tensorflow_graphics.geometry.transformation'] = tfg.geometry.transformation
    dep = os.path.join(REFERENCE_ROOT, 'dependencies', 'manipulation_tasks')
    if os.path.isdir(dep) and dep not in sys.path:
        sys.path.append(dep)          # the reference's own Affine (numpy / scipy only)
    for opt in ('ftfy', 'matplotlib', 'matplotlib.pyplot', 'PIL', 'PIL.Image'):
        try:
            importlib.import_module(opt)
        except Exception:
            sys.modules[opt] = types.ModuleType(opt)
            if '.' in opt:
                setattr(sys.modules[opt.split('.')[0]], opt.split('.')[1], sys.modules[opt])
    return tf


def load_reference():
    """import the reference's ``src`` package as ``refsrc`` and return
    (nerf_utils, layers, model_v0) modules.  Needs the reference checkout (build container only)."""
    install()
    src = os.path.join(REFERENCE_ROOT, 'src')
    if not os.path.isdir(src):
        raise FileNotFoundError(f'{src} not found (the reference checkout exists only in the build container)')
    for name, path in (('refsrc', src), ('refsrc.lib', os.path.join(src, 'lib'))):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__path__ = [path]
            sys.modules[name] = m
    nu = importlib.import_module('refsrc.lib.mvnerf.nerf_utils')
    ly = importlib.import_module('refsrc.lib.mvnerf.layers')
    mv = importlib.import_module('refsrc.lib.mvnerf.model_v0')
    return nu, ly, mv


def load_reference_query(which: str = 'delta_ngf'):
    """the reference's pose-query model module and its GraspReadout layers module:
    'delta_ngf' -> (delta_ngf.model, delta_ngf.layers), 'grasp_mvnerf' -> (grasp_mvnerf.model, grasp_mvnerf.layers)"""
    load_reference()
    return (importlib.import_module(f'refsrc.lib.{which}.model'), importlib.import_module(f'refsrc.lib.{which}.layers'))
