"""cn_grasp_head_bwd_static: the allocation-free variant of cn_grasp_head_bwd (scratch in a file-scope ``static
__device__`` array of the library instead of a caller-provided ``dpc_ws``; no counterpart in the reference, see
include/clipnerf_b200.h).  Called through the C ABI and compared with the existing entry point on the same inputs:
the kernels are the same, so the bar is the one the repo uses for two launches of one kernel -- identical bits.
Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _head_case(cuda, pack, b, n_poses, seed, elu=True):
    """forward of the head on seeded taps -> (gy, HeadState) for the two backward entry points"""
    from clipnerf_b200 import ops
    g = torch.Generator().manual_seed(seed)
    n_pad = (n_poses * 42 + 127) // 128 * 128
    taps = (torch.randn(4, b * n_pad, 128, generator=g) * 0.7).to(cuda)
    gy = torch.randn(b, n_poses, generator=g).to(cuda)
    _, st = ops.grasp_head_fwd(taps, b, n_pad, n_poses, pack, elu, True)
    return gy, st


def _bwd_static(gy, st, out):
    """the new entry point through the C ABI; every buffer is the caller's, ``out`` included"""
    from clipnerf_b200 import _abi, ops
    _abi.call('cn_grasp_head_bwd_static', ops._ptr(gy), st.n_tiles, st.b, st.n_pad, st.np, ops._ptr(st.pack), st.elu,
              ops._ptr(st.vslabs), ops._ptr(st.acat), ops._ptr(st.saved), ops._ptr(out), ops._stream())
    return out


def _pack(cuda):
    from clipnerf_b200 import _abi, ops
    params = torch.randn(_abi.HEAD_PARAMS, generator=torch.Generator().manual_seed(7)) * 0.05
    return ops.grasp_head_pack(params.to(cuda))


def test_static_scratch_backward_equals_the_existing_entry_point_on_calls_in_a_row(cuda):
    from clipnerf_b200 import ops
    pack = _pack(cuda)
    # two different shapes back to back (the second sees the first call's contents in the shared array), then the
    # first again, then the largest accepted size (1024 poses = 8 pose tiles)
    cases = [_head_case(cuda, pack, 2, 200, 1), _head_case(cuda, pack, 1, 3, 2, elu=False)]
    cases += [cases[0], _head_case(cuda, pack, 8, 128, 3)]
    outs = [torch.empty(4, st.n_tiles * ops.ACT_TILE_BYTES, dtype=torch.uint8, device=cuda) for _, st in cases]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for (gy, st), out in zip(cases, outs):
        _bwd_static(gy, st, out)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before, 'the static-scratch call went through the torch allocator'
    for (gy, st), out in zip(cases, outs):
        ref = ops.grasp_head_bwd(gy, st)
        assert torch.isfinite(out.view(torch.bfloat16).float()).all()
        assert torch.equal(out, ref), f'b={st.b} np={st.np}: differs from cn_grasp_head_bwd'
    assert not torch.equal(outs[0], torch.zeros_like(outs[0]))


def test_static_scratch_backward_rejects_sizes_beyond_its_maximum(cuda):
    from clipnerf_b200 import _abi, ops
    assert _abi.HEAD_BWD_STATIC_MAX_POSES == 1024
    pack = _pack(cuda)
    gy, st = _head_case(cuda, pack, 5, 205, 4)                   # 1025 poses
    out = torch.full((4, st.n_tiles * ops.ACT_TILE_BYTES), 0x5a, dtype=torch.uint8, device=cuda)
    with pytest.raises(RuntimeError, match='CN_HEAD_BWD_STATIC_MAX_POSES'):
        _bwd_static(gy, st, out)
    torch.cuda.synchronize()
    assert bool((out == 0x5a).all()), 'a rejected call must not launch anything'
    # the existing entry point takes the same case, and the static one still works afterwards
    ops.grasp_head_bwd(gy, st)
    gy, st = _head_case(cuda, pack, 1, 7, 5)
    out = torch.empty(4, st.n_tiles * ops.ACT_TILE_BYTES, dtype=torch.uint8, device=cuda)
    assert torch.equal(_bwd_static(gy, st, out), ops.grasp_head_bwd(gy, st))
