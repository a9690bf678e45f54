"""Process grid and block decomposition (Python view of csrc/core/decomp.*).

Parity: ``choose_process_grid`` S4:82-86 and ``decompose_2d`` S4:97-133 of the reference; the answers are
identical for every rank count (tests/test_decomp.py checks P = 1..32 against a literal restatement).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

from ..ops import _native


@dataclass(frozen=True)
class Block:
    rank: int
    px: int
    py: int
    i0: int
    j0: int
    nx: int
    ny: int
    pitch: int
    nbr: tuple  # x-, x+, y-, y+ neighbour ranks or -1

    @property
    def i_range(self):
        return (self.i0, self.i0 + self.nx - 1)

    @property
    def j_range(self):
        return (self.j0, self.j0 + self.ny - 1)


def choose_process_grid(nranks: int) -> tuple[int, int]:
    """``(Px, Py)`` with ``Px <= Py``, ``Px`` the largest divisor of ``nranks`` not above its square root."""
    px, py = C.c_int(), C.c_int()
    _native.lib().pb_choose_proc_grid(int(nranks), C.byref(px), C.byref(py))
    return px.value, py.value


def nvswitch_process_grid(nranks: int, M: int) -> tuple[int, int]:
    """Process grid for GPUs behind an NVSwitch (every peer equally close, halo bytes negligible next to HBM
    traffic): split along x only (``nranks x 1``).  Every shared edge is then a contiguous row of the arrays
    (coalesced peer stores, no strided column edges) and the edge tiles are the first tiles of every phase, so
    the release of the halo stores overlaps with the interior.  Falls back to the reference's near-square grid
    when the slabs would get thinner than one 32-row tile."""
    if nranks > 1 and (M - 1) // nranks >= 32:
        return nranks, 1
    return choose_process_grid(nranks)


def parse_grid(spec, nranks: int, M: int | None = None) -> tuple[int, int]:
    """``None`` / ``"reference"`` -> the reference's choice; ``"auto"`` -> :func:`nvswitch_process_grid`;
    ``"2x4"`` or ``(2, 4)`` -> explicit (must multiply to ``nranks``)."""
    if spec is None or spec == "reference":
        return choose_process_grid(nranks)
    if spec == "auto":
        return nvswitch_process_grid(nranks, M if M is not None else 1 << 30)
    if isinstance(spec, str):
        a, b = spec.lower().split("x")
        spec = (int(a), int(b))
    px, py = int(spec[0]), int(spec[1])
    if px * py != nranks:
        raise ValueError(f"process grid {px}x{py} does not match {nranks} ranks")
    return px, py


def local_block(M: int, N: int, Px: int, Py: int, rank: int) -> Block:
    b = _native.Block()
    _native.check(_native.lib().pb_local_block(M, N, Px, Py, rank, C.byref(b)), "pb_local_block")
    return Block(b.rank, b.px, b.py, b.i0, b.j0, b.nx, b.ny, b.pitch, tuple(b.nbr))


def native_block(blk: Block) -> "_native.Block":
    b = _native.Block()
    b.rank, b.px, b.py, b.i0, b.j0, b.nx, b.ny, b.pitch = (blk.rank, blk.px, blk.py, blk.i0, blk.j0, blk.nx,
                                                         blk.ny, blk.pitch)
    for d in range(4):
        b.nbr[d] = blk.nbr[d]
    return b


def col_offset() -> int:
    return _native.lib().pb_col_offset()
