"""Pins the tcgen05 shared-memory descriptor encoding the MLP kernels rely on: one 128x128
accumulator tile from caller-built operand images, compared with a CPU matmul."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _bf16_mat(rows, cols, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(rows, cols, generator=g).to(torch.bfloat16).to(torch.float32)


def test_kmajor_sw128(cuda):
    from clipnerf_b200 import layout, ops
    a = _bf16_mat(128, 64, 1)           # A [M=128, K=64]
    b = _bf16_mat(128, 64, 2)           # B [N=128, K=64]
    d = ops.probe_umma(layout.matrix_to_tiles(a).to(cuda), layout.matrix_to_tiles(b).to(cuda), 4, False, False,
                       16, 1024, 32, 16, 1024, 32)
    ref = a @ b.t()
    assert rel_err(d, ref) < 1e-5, rel_err(d, ref)


def test_kmajor_two_slabs(cuda):
    """K = 128: the MLP kernels step +32 B inside a slab and +16 KiB across slabs."""
    from clipnerf_b200 import layout, ops
    a = _bf16_mat(128, 128, 3)
    b = _bf16_mat(128, 128, 4)
    ta, tb = layout.matrix_to_tiles(a).to(cuda), layout.matrix_to_tiles(b).to(cuda)
    acc = torch.zeros(128, 128)
    for s in range(2):
        d = ops.probe_umma(ta[s * 16384:(s + 1) * 16384].contiguous(), tb[s * 16384:(s + 1) * 16384].contiguous(), 4,
                           False, False, 16, 1024, 32, 16, 1024, 32)
        acc += d.cpu()
    assert rel_err(acc, a @ b.t()) < 1e-5


def test_mnmajor_sw128(cuda):
    """wgrad form: D[m, n] = sum_k A[k, m] B[k, n] with both tiles stored [k][m|n] (rows = K)."""
    from clipnerf_b200 import layout, ops
    a = _bf16_mat(128, 128, 5)          # [K=128 rows, M=128]
    b = _bf16_mat(128, 128, 6)          # [K=128 rows, N=128]
    d = ops.probe_umma(layout.matrix_to_tiles(a).to(cuda), layout.matrix_to_tiles(b).to(cuda), 8, True, True,
                       16384, 1024, 2048, 16384, 1024, 2048)
    ref = a.t() @ b
    assert rel_err(d, ref) < 1e-5, rel_err(d, ref)
