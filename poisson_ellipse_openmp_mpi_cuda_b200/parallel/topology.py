"""Process grid from the GPUs' peer-to-peer topology.

The reference picks its Px x Py grid from the rank count alone (``choose_process_grid``, S4:82-86); here the
grid and the assignment of blocks to devices follow what CUDA reports about the links between the devices
(``cudaDeviceGetP2PAttribute``).  The decision itself lives in C++ (csrc/core/decomp.cpp,
``choose_grid_from_topology``) and is shared with the CLI; this module probes the matrix and calls it.
"""
from __future__ import annotations

import ctypes as C

from ..ops import _native

KINDS = {0: "uniform", 1: "islands", 2: "fallback"}


def probe_p2p(devices) -> dict:
    """``{"perf", "access", "atomics"}``: n x n lists for the given device ordinals (needs CUDA)."""
    devs = list(devices)
    n = len(devs)
    arr = (C.c_int * n)(*devs)
    out = [(C.c_int * (n * n))() for _ in range(3)]
    _native.check(_native.lib().pb_p2p_matrix(arr, n, out[0], out[1], out[2]), "pb_p2p_matrix")
    return {k: [[int(o[a * n + b]) for b in range(n)] for a in range(n)] for k, o in zip(("perf", "access", "atomics"), out)}


def grid_from_topology(perf, M: int) -> tuple[int, int, list[int], str]:
    """``(Px, Py, order, kind)`` for a performance-rank matrix (lower = better link; diagonal ignored).

    ``order[rank]`` is the index of the device that runs block ``rank``.  Uniform fabric (NVSwitch): ``n x 1`` slabs
    in device order; equal islands of well-connected devices: ``islands x members`` with the x-cuts between the
    islands; anything else: the reference's grid in device order.
    """
    n = len(perf)
    flat = (C.c_int * (n * n))(*[int(perf[a][b]) for a in range(n) for b in range(n)])
    px, py = C.c_int(), C.c_int()
    order = (C.c_int * n)()
    kind = _native.lib().pb_choose_grid_topology(n, flat, int(M), C.byref(px), C.byref(py), order)
    if kind < 0:
        raise RuntimeError(_native.last_error())
    return px.value, py.value, [int(v) for v in order], KINDS[kind]
