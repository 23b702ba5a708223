"""bf16 tile images (see ``include/clipnerf_b200.h``): conversions between the swizzled slab layout
the tcgen05 kernels consume and plain row-major matrices.  Used by tests / diagnostics and by
``render`` for nothing on the hot path (the kernels read and write tile images directly)."""
from __future__ import annotations

import functools

import torch

from ._abi import SLAB_BYTES, TILE_ROWS


@functools.lru_cache(maxsize=None)
def _elem_index(ncols: int) -> torch.Tensor:
    """[128, ncols] int64: bf16-element offset of (row, col) inside one tile image."""
    r = torch.arange(TILE_ROWS).view(-1, 1)
    c = torch.arange(ncols).view(1, -1)
    byte = (c // 64) * SLAB_BYTES + r * 128 + ((((c // 8) % 8) ^ (r % 8)) * 16) + (c % 8) * 2
    return (byte // 2).to(torch.int64)


@functools.lru_cache(maxsize=None)
def _elem_index_on(ncols: int, device: str) -> torch.Tensor:
    """the index table on a device, copied once (a host -> device copy cannot be recorded into a CUDA graph)"""
    return _elem_index(ncols).to(torch.device(device)).view(-1)


def x_col_to_ref(j: int) -> int:
    """internal column of the MLP input X -> reference column (layers.py:362), -1 = zero pad."""
    if j < 256:
        return 123 + j
    if j < 316:
        return j - 256
    if j < 376:
        return 60 + (j - 316)
    if j < 379:
        return 120 + (j - 376)
    return -1


def tiles_to_matrix(buf: torch.Tensor, ncols: int) -> torch.Tensor:
    """uint8 tile images -> [n_tiles*128, ncols] float32."""
    elems = ncols * TILE_ROWS
    t = buf.view(torch.bfloat16).view(-1, elems)
    idx = _elem_index_on(ncols, str(buf.device))
    return t[:, idx].view(-1, ncols).to(torch.float32)


def matrix_to_tiles(mat: torch.Tensor) -> torch.Tensor:
    """[n_tiles*128, ncols] float -> uint8 tile images (bf16, round-to-nearest-even)."""
    rows, ncols = mat.shape
    assert rows % TILE_ROWS == 0 and ncols % 64 == 0
    nt = rows // TILE_ROWS
    out = torch.zeros(nt, ncols * TILE_ROWS, dtype=torch.bfloat16, device=mat.device)
    idx = _elem_index_on(ncols, str(mat.device))
    out[:, idx] = mat.to(torch.bfloat16).view(nt, -1)
    return out.view(torch.uint8).view(-1)


def x_matrix_to_ref(xm: torch.Tensor) -> torch.Tensor:
    """[rows, 384] internal column order -> [rows, 379] reference order."""
    ref = torch.empty(xm.shape[0], 379, dtype=xm.dtype, device=xm.device)
    cols = torch.tensor([x_col_to_ref(j) for j in range(379)], device=xm.device)
    ref[:, cols] = xm[:, :379]
    return ref
