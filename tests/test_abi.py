"""The C-ABI shared library loads without a GPU and exports every symbol include/*.h declares;
the host-side binding lists exactly the same entry points."""
import ctypes
import glob
import os
import re

from conftest import ROOT


def _declared():
    names = []
    for h in glob.glob(os.path.join(ROOT, 'include', '*.h')):
        src = open(h).read()
        src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
        names += re.findall(r'\b(cn_[a-z0-9_]+)\s*\(', src)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    from clipnerf_b200 import _abi
    assert os.path.exists(_abi.LIB_PATH), 'run __graft_entry__.build() first'
    lib = ctypes.CDLL(_abi.LIB_PATH)
    declared = _declared()
    assert len(declared) >= 19
    for name in declared:
        assert hasattr(lib, name), f'{name} declared in include/ but not exported'


def test_binding_matches_header():
    from clipnerf_b200 import _abi
    assert sorted(_abi.PROTOTYPES) == _declared()
    assert _abi.load().cn_abi_version() == _abi.ABI_VERSION


def test_argument_errors_do_not_need_a_gpu():
    """shape / pointer validation happens on the host before any launch: rc < 0 and a message."""
    from clipnerf_b200 import _abi
    lib = _abi.load()
    rc = lib.cn_composite_fwd(None, None, 4, 64, None, None, None, None)
    assert rc < 0 and b'null pointer' in lib.cn_last_error()
    one = ctypes.c_void_p(16)
    rc = lib.cn_composite_fwd(one, one, 4, 48, one, one, one, None)
    assert rc < 0 and b'S must be 64 or 128' in lib.cn_last_error()
    rc = lib.cn_mlp_block_bwd(one, 9, one, 6, 0, 1, one, 1, 0, one, 1.0, one, None)
    assert rc < 0 and b'blk' in lib.cn_last_error()
