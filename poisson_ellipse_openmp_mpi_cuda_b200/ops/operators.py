"""Stand-alone device operators on blocks in the padded layout (thin wrappers over the C ABI)."""
from __future__ import annotations

import torch

from . import _native
from .gpu import _stream_ptr


def apply_A_dot(M: int, N: int, p: torch.Tensor, a: torch.Tensor, b: torch.Tensor, nx: int | None = None,
                ny: int | None = None, out: torch.Tensor | None = None):
    """``Ap = A p`` and ``(Ap, p) = h1 h2 sum(Ap * p)`` on one block.

    ``p, a, b``: float64 CUDA tensors of shape ``(nx+2, pitch)`` in the padded layout (ring included, column
    ``lj`` at ``lj + col_offset``).  Returns ``(Ap, dot)`` with ``dot`` a one-element device tensor; nothing is
    synchronised or copied to the host.  Replaces ``apply_A_kernel`` + ``dot_gpu`` (S4:507-536, S4:771-786).
    """
    lib = _native.lib()
    co = lib.pb_col_offset()
    nx = p.shape[0] - 2 if nx is None else nx
    ny = (N - 1) if ny is None else ny
    pitch = p.shape[1]
    assert p.is_cuda and p.dtype == torch.float64 and p.is_contiguous() and a.shape == p.shape == b.shape
    assert pitch >= ny + 2 + co
    Ap = torch.zeros_like(p) if out is None else out
    dot = torch.zeros(1, dtype=torch.float64, device=p.device)
    with torch.cuda.device(p.device):
        _native.check(lib.pb_apply_A_dot(M, N, nx, ny, pitch, p.data_ptr(), a.data_ptr(), b.data_ptr(), Ap.data_ptr(),
                                         dot.data_ptr(), _stream_ptr(None)), "pb_apply_A_dot")
    return Ap, dot
