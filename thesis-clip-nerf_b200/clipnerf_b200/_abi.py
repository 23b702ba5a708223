"""ctypes binding of ``libclipnerf_b200.so`` (the C ABI declared in ``include/clipnerf_b200.h``).

The library is built in-tree by ``__graft_entry__.build()`` / ``csrc/Makefile``.  There is no
CPU fallback: if the shared object is missing every op raises."""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_double, c_float, c_int32, c_int64, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), 'lib', 'libclipnerf_b200.so')

ABI_VERSION = 11
TILE_ROWS = 128
SLAB_BYTES = 16384
ACT_TILE_BYTES = 32768
X_TILE_BYTES = 98304
EMB_PARAMS = 246784
NET_PARAMS = 247300
PACKED_BYTES = 2 * 491520 + 16384
HEAD_PARAMS = 582337
HEAD_PACK_BYTES = 2329600
HEAD_OFFSETS = 42
HEAD_BWD_STATIC_MAX_POSES = 1024   # CN_HEAD_BWD_STATIC_MAX_POSES: poses cn_grasp_head_bwd_static's built-in scratch holds


class CnScene(ctypes.Structure):
    _fields_ = [('images', c_void_p), ('features', c_void_p), ('intrinsics', c_void_p), ('extr_inv', c_void_p),
                ('b', c_int32), ('nv', c_int32), ('H', c_int32), ('W', c_int32), ('feat_up', c_int32)]


_P = c_void_p
# name -> argtypes (restype is int unless listed in _RESTYPES)
PROTOTYPES = {
    'cn_abi_version': [],
    'cn_last_error': [],
    'cn_sample_coarse': [_P, _P, c_int64, c_int32, c_double, c_double, _P],
    'cn_sample_pdf_fwd': [_P, _P, _P, c_int64, _P, _P, _P, _P],
    'cn_sample_pdf': [_P, _P, _P, c_int64, _P, _P, _P],
    'cn_sample_pdf_bwd': [_P, _P, _P, _P, _P, c_int64, _P, _P],
    'cn_gather_ws_bytes': [c_int32, c_int32, c_int32, c_int32, c_int32, c_int32],
    'cn_project_bin': [ctypes.POINTER(CnScene), _P, _P, _P, c_int32, c_int32, _P, c_size_t, _P],
    'cn_project_gather_fwd': [ctypes.POINTER(CnScene), _P, _P, _P, c_int32, c_int32, _P, _P, c_size_t, _P],
    'cn_project_gather_debug': [ctypes.POINTER(CnScene), _P, _P, _P, c_int32, c_int32, _P, _P, _P, _P, _P],
    'cn_interpolate_bilinear': [_P, _P, c_int32, c_int32, c_int32, c_int32, c_int64, _P, _P],
    'cn_position_encoding': [_P, c_int64, c_int32, c_int32, c_float, _P, _P],
    'cn_project_gather_bwd': [ctypes.POINTER(CnScene), _P, _P, _P, c_int32, c_int32, _P, _P, c_size_t, _P, _P, _P],
    'cn_project_gather_bwd_rays': [ctypes.POINTER(CnScene), _P, _P, _P, c_int32, c_int32, _P, _P, c_size_t, _P, _P, _P,
                                   _P],
    'cn_feature_grad': [ctypes.POINTER(CnScene), c_int32, _P, _P, _P, _P, _P, _P, c_int32, _P],
    'cn_generate_rays': [_P, _P, c_int32, c_int32, c_int64, c_int64, _P, _P, _P],
    'cn_quantize_render': [_P, _P, c_int64, _P, _P, _P, _P],
    'cn_images_u8_to_f32': [_P, c_int64, c_int32, _P, _P],
    'cn_composite_fwd': [_P, _P, c_int64, c_int32, _P, _P, _P, _P],
    'cn_composite_bwd': [_P, _P, c_int64, c_int32, _P, _P, _P, _P, _P, _P],
    'cn_pack_weights': [_P, _P, _P],
    'cn_mlp_feature_fwd': [_P, _P, c_int64, _P, _P, _P],
    'cn_mlp_fusion_fwd': [_P, _P, c_int32, c_int32, c_int64, _P, _P, _P],
    'cn_mlp_chain_fwd': [_P, _P, c_int64, _P, _P, _P, _P, c_int32, c_int32, _P, _P, _P],
    'cn_mlp_fusion_fwd_taps': [_P, _P, c_int32, c_int32, c_int64, _P, _P, _P, _P],
    'cn_tile_add_rows': [_P, _P, c_int64, c_int32, _P],
    'cn_mlp_fusion_fwd_tap_tiles': [_P, _P, c_int32, c_int32, c_int64, _P, _P, _P, _P],
    'cn_tile_add_tiles': [_P, _P, c_int64, _P],
    'cn_query_tangent_x': [ctypes.POINTER(CnScene), _P, _P, _P, _P, c_int32, _P, _P],
    'cn_mlp_feature_fwd_ex': [_P, _P, c_int64, _P, _P, c_int32, _P, _P],
    'cn_mlp_fusion_fwd_taps_ex': [_P, _P, c_int32, c_int32, c_int64, _P, _P, _P, c_int32, c_int32, _P, _P],
    'cn_pose_points_fwd': [_P, _P, c_int64, c_int32, _P, _P, _P],
    'cn_pose_points_bwd': [_P, _P, _P, c_int64, c_int32, _P, _P],
    'cn_grasp_head_pack': [_P, _P, _P],
    'cn_grasp_head_fwd': [_P, c_int64, c_int32, c_int32, c_int32, _P, c_int32, _P, _P, _P, _P, _P],
    'cn_grasp_head_bwd': [_P, c_int64, c_int32, c_int32, c_int32, _P, c_int32, _P, _P, _P, _P, _P, _P],
    'cn_grasp_head_bwd_static': [_P, c_int64, c_int32, c_int32, c_int32, _P, c_int32, _P, _P, _P, _P, _P],
    'cn_mlp_readout_bwd': [_P, _P, _P, c_int32, c_int32, c_int64, _P, _P, _P],
    'cn_mlp_block_bwd': [_P, c_int32, _P, c_int32, c_int32, c_int64, _P, c_int32, c_int64, _P, c_float, _P, _P],
    'cn_mlp_stage_bwd_ws_bytes': [],
    'cn_mlp_stage_bwd': [_P, c_int32, _P, c_int32, c_int64, _P, c_int32, c_int64, _P, _P, _P, c_float, _P, _P, c_size_t, _P],
    'cn_mlp_mean_bwd': [_P, c_int32, c_int32, c_int64, _P, _P],
    'cn_mlp_l0_dgrad': [_P, _P, c_int64, _P, c_float, _P],
    'cn_mlp_l0_wgrad': [_P, _P, c_int64, c_float, _P, _P],
    'cn_adam_step': [_P, _P, _P, _P, c_int64, c_int64, c_int64, _P, c_double, c_double, c_double, c_float, c_float,
                     c_double, c_double, c_double, _P],
    'cn_probe_umma': [_P, c_int32, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32,
                      c_int32, c_int32, _P, _P],
    'cn_probe_tile_reduce': [_P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, _P],
}
_RESTYPES = {'cn_last_error': c_char_p, 'cn_gather_ws_bytes': c_size_t, 'cn_mlp_stage_bwd_ws_bytes': c_size_t}

_lib = None


def load():
    """Load the shared object (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f'{LIB_PATH} is missing: run `python -c "import __graft_entry__ as g; g.build()"` '
                               f'(there is no CPU fallback)')
        lib = ctypes.CDLL(LIB_PATH)
        for name, argtypes in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, ctypes.c_int)
        if lib.cn_abi_version() != ABI_VERSION:
            raise RuntimeError(f'ABI version mismatch: library {lib.cn_abi_version()} != binding {ABI_VERSION}')
        _lib = lib
    return _lib


class CnError(RuntimeError):
    pass


# kernels launched by each entry point (for launch accounting in bench.py)
KERNELS_PER_CALL = {name: 1 for name in PROTOTYPES}
KERNELS_PER_CALL.update({'cn_abi_version': 0, 'cn_last_error': 0, 'cn_gather_ws_bytes': 0, 'cn_mlp_stage_bwd_ws_bytes': 0,
                         'cn_project_bin': 4, 'cn_project_gather_fwd': 5, 'cn_feature_grad': 2,
                         'cn_project_gather_bwd_rays': 2, 'cn_quantize_render': 2, 'cn_adam_step': 2,
                         'cn_grasp_head_fwd': 2, 'cn_grasp_head_bwd': 2, 'cn_grasp_head_bwd_static': 2})

LAUNCHES = 0          # running count of kernel launches issued through the ABI
TIMER = None          # optional callable(name) -> context manager, set by bench.py to time stages


def call(name: str, *args, extra_launches=None):
    global LAUNCHES
    lib = load()
    if TIMER is not None:
        with TIMER(name):
            rc = getattr(lib, name)(*args)
    else:
        rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.cn_last_error().decode('utf-8', 'replace')
        raise CnError(f'{name} failed (rc={rc}): {msg}')
    LAUNCHES += KERNELS_PER_CALL[name] if extra_launches is None else extra_launches
