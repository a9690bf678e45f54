"""GPU solver objects: a thin, typed wrapper around ``pb_solver`` (C ABI) that hands out torch views.

The device arrays are owned by the native arena (one ``cudaMalloc`` per rank, exportable through CUDA
IPC); torch tensors returned here alias that memory through ``__cuda_array_interface__``.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from . import _native

NORMS = {"weighted": 0, "unweighted": 1}
DIVS = {"recip": 0, "true": 1, "markstein": 2}
METHODS = {"persistent": 0, "step_fused": 1, "step_classic": 2, "persistent_ldg": 3, "resident": 4, "persistent_tma": 5}
STATUS = {0: "max_iter", 1: "converged", 2: "breakdown", 3: "running", 4: "watchdog", 5: "tma_timeout"}
ARRAYS = {"w": 0, "r": 1, "p0": 2, "p1": 3, "Ap": 4, "a": 5, "b": 6, "D": 7}


class _DevMem:
    """Minimal ``__cuda_array_interface__`` carrier for memory owned by the native library."""

    def __init__(self, ptr: int, shape, typestr: str, owner):
        self.__cuda_array_interface__ = {
            "shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3, "strides": None,
        }
        self._owner = owner  # keeps the solver alive while views exist


def _stream_ptr(stream) -> int:
    if stream is None:
        stream = torch.cuda.current_stream()
    return int(stream.cuda_stream)


@dataclass
class BlockInfo:
    rank: int
    px: int
    py: int
    i0: int
    j0: int
    nx: int
    ny: int
    pitch: int
    nbr: tuple


class GpuSolver:
    """Owns the arenas of ``nlocal`` consecutive ranks of a ``Px x Py`` grid on one device."""

    def __init__(self, M: int, N: int, Px: int = 1, Py: int = 1, first_rank: int = 0, nlocal: int | None = None,
                 device: int | None = None):
        if not torch.cuda.is_available():
            raise RuntimeError("GpuSolver needs a CUDA device (torch.cuda.is_available() is False)")
        self.lib = _native.lib()
        self.M, self.N, self.Px, self.Py = M, N, Px, Py
        self.first_rank = first_rank
        self.nlocal = Px * Py if nlocal is None else nlocal
        self.device = torch.cuda.current_device() if device is None else device
        torch.cuda.init()
        self._h = self.lib.pb_solver_create(M, N, Px, Py, first_rank, self.nlocal, self.device)
        if not self._h:
            raise RuntimeError(f"pb_solver_create failed: {_native.last_error()}")
        self._finalized = False
        self.col_off = self.lib.pb_col_offset()

    # -- lifetime -------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self.lib.pb_solver_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- peers ----------------------------------------------------------------------------------
    def ipc_handle(self, local_idx: int = 0) -> bytes:
        buf = C.create_string_buffer(64)
        _native.check(self.lib.pb_solver_ipc_handle(self._h, local_idx, buf), "pb_solver_ipc_handle")
        return buf.raw

    def open_peer(self, rank: int, handle: bytes):
        _native.check(self.lib.pb_solver_open_peer(self._h, rank, C.create_string_buffer(handle, 64)),
                      "pb_solver_open_peer")

    def finalize(self):
        _native.check(self.lib.pb_solver_finalize(self._h), "pb_solver_finalize")
        self._finalized = True
        return self

    # -- info -----------------------------------------------------------------------------------
    def block(self, local_idx: int = 0) -> BlockInfo:
        b = _native.Block()
        _native.check(self.lib.pb_solver_block(self._h, local_idx, C.byref(b)), "pb_solver_block")
        return BlockInfo(b.rank, b.px, b.py, b.i0, b.j0, b.nx, b.ny, b.pitch, tuple(b.nbr))

    def array(self, name: str, local_idx: int = 0) -> torch.Tensor:
        """Padded ``(nx+2, pitch)`` float64 view of a device array (ring at row/col 0 and nx+1/ny+1;
        column ``lj`` lives at ``lj + col_off``)."""
        blk = self.block(local_idx)
        ptr = self.lib.pb_solver_array(self._h, local_idx, ARRAYS[name])
        if not ptr:
            raise RuntimeError(_native.last_error())
        return torch.as_tensor(_DevMem(ptr, (blk.nx + 2, blk.pitch), "<f8", self), device=f"cuda:{self.device}")

    def interior(self, name: str, local_idx: int = 0) -> torch.Tensor:
        blk = self.block(local_idx)
        return self.array(name, local_idx)[1:blk.nx + 1, self.col_off + 1:self.col_off + 1 + blk.ny]

    def control_red(self, local_idx: int = 0) -> torch.Tensor:
        ptr = self.lib.pb_solver_control_red(self._h, local_idx)
        return torch.as_tensor(_DevMem(ptr, (4,), "<f8", self), device=f"cuda:{self.device}")

    def control_err(self, local_idx: int = 0) -> torch.Tensor:
        ptr = self.lib.pb_solver_control_err(self._h, local_idx)
        return torch.as_tensor(_DevMem(ptr, (5,), "<f8", self), device=f"cuda:{self.device}")

    def info(self) -> dict:
        ctas, launches, h2d, d2h = C.c_int(), C.c_longlong(), C.c_longlong(), C.c_longlong()
        self.lib.pb_solver_info(self._h, C.byref(ctas), C.byref(launches), C.byref(h2d), C.byref(d2h))
        return {"ctas_per_rank": ctas.value, "launches": launches.value, "h2d_bytes": h2d.value,
                "d2h_bytes": d2h.value}

    def cta_times(self, local_idx: int = 0) -> torch.Tensor:
        """Per-CTA device clocks of the last persistent solve: ``[ctas, 4]`` seconds in phase AB, phase C and
        waiting in reductions 1 and 2 (a CTA that waits long in a reduction finished its phase early)."""
        n = self.info()["ctas_per_rank"]
        out = torch.zeros((max(n, 1), 4), dtype=torch.float64)
        got = self.lib.pb_solver_cta_times(self._h, local_idx, C.cast(out.data_ptr(), C.POINTER(C.c_double)), n)
        _native.check(min(got, 0), "pb_solver_cta_times")
        return out[:max(got, 0)]

    def cta_report(self, local_idx: int = 0) -> torch.Tensor:
        """Raw per-CTA report words ``[ctas, 8]`` (see ``pb_solver_cta_report``)."""
        n = self.info()["ctas_per_rank"]
        out = torch.zeros((max(n, 1), 8), dtype=torch.float64)
        got = self.lib.pb_solver_cta_report(self._h, local_idx, C.cast(out.data_ptr(), C.POINTER(C.c_double)), n)
        _native.check(min(got, 0), "pb_solver_cta_report")
        return out[:max(got, 0)]

    def tile_counts(self) -> dict:
        """Tiles of the last persistent (TMA) solve by bytes moved per node and iteration (counted by the kernel)."""
        out = (C.c_longlong * 3)()
        _native.check(self.lib.pb_solver_tile_counts(self._h, out), "pb_solver_tile_counts")
        return {"recompute_64B": out[0], "stored_80B": out[1], "mixed_112B": out[2]}

    def recomputes_ap(self, ap_mode: int = 0) -> bool:
        """Whether a solve with this ``ap_mode`` recomputes Ap in phase C (0 = decided by working-set size vs L2)."""
        return bool(self.lib.pb_solver_ap_recompute(self._h, int(ap_mode)))

    # -- solve ----------------------------------------------------------------------------------
    def solve(self, method: str = "persistent", norm: str = "weighted", delta: float = 1e-6, max_iter: int = -1,
              exact_div: bool | None = None, ctas_per_sm: int = 0, error_report: bool = True,
              watchdog_s: float = 0.0, tile_rows: int = 0, div: str = "markstein", l2_hints: bool = False,
              full_coefficients: bool = False, ap_mode: int = 0, balance: bool = True, ab_rows: bool = True,
              stream=None) -> dict:
        """``div``: how the stencil divides by h1/h2 -- "true" (IEEE division), "markstein" (three-instruction
        sequence with the same bits, default) or "recip" (multiply by the reciprocal).  ``exact_div`` is the
        older boolean spelling: True -> "true", False -> "recip".

        TMA kernel only (see ``pb_solve_options`` in include/poisson_b200.h): ``tile_rows`` 0 = auto, 16 / 32 / 64;
        ``ap_mode`` 0 = auto, 1 = Ap stored, 2 = recomputed in phase C; ``balance`` False = cyclic tile -> CTA map
        instead of the cost-balanced one; ``ab_rows`` False = phase AB on the tile path for every tile instead of
        warp-private rows on full constant-coefficient tiles; ``full_coefficients`` streams a, b, D instead of the
        class bytes; ``l2_hints`` puts eviction priorities on the TMA loads.  All of them leave every value equal to
        rounding; ``balance`` and ``ab_rows`` change the order of the partial sums."""
        if not self._finalized:
            self.finalize()
        if exact_div is not None:
            div = "true" if exact_div else "recip"
        opt = _native.SolveOptions(METHODS[method], NORMS[norm], delta, max_iter, DIVS[div], ctas_per_sm,
                                   int(error_report), tile_rows, watchdog_s, int(l2_hints), int(full_coefficients),
                                   int(ap_mode), int(not balance), int(not ab_rows))
        st = _native.Stats()
        with torch.cuda.device(self.device):
            _native.check(self.lib.pb_solver_solve(self._h, C.byref(opt), _stream_ptr(stream), C.byref(st)),
                          "pb_solver_solve")
        out = st.as_dict()
        out["status_name"] = STATUS.get(st.status, "?")
        out.update(self.info())
        out["kernel"] = (self.lib.pb_solver_last_kernel(self._h) or b"").decode()
        if st.status in (4, 5):
            raise RuntimeError(f"device-side {STATUS[st.status]}: the solve was abandoned (debug: {out})")
        return out

    def solution_host(self) -> torch.Tensor:
        """Dense ``(M+1, N+1)`` host field; only nodes owned by local ranks are filled."""
        full = torch.zeros((self.M + 1, self.N + 1), dtype=torch.float64).pin_memory()
        _native.check(self.lib.pb_solver_copy_solution(self._h, full.data_ptr(), _stream_ptr(None)),
                      "pb_solver_copy_solution")
        return full

    def gather_block(self, local_idx: int = 0) -> torch.Tensor:
        """Dense ``(nx, ny)`` device copy of the owned nodes of ``w`` (native gather kernel)."""
        blk = self.block(local_idx)
        out = torch.empty((blk.nx, blk.ny), dtype=torch.float64, device=f"cuda:{self.device}")
        _native.check(self.lib.pb_solver_gather_block(self._h, local_idx, out.data_ptr(), blk.ny, _stream_ptr(None)),
                      "pb_solver_gather_block")
        return out

    def xgpu_bench(self, iters: int, mode: int = 0, edge: int = 0, ctas: int = 0) -> tuple[float, float]:
        """Communication micro-benchmark (collective): ns per round on this rank, checksum."""
        out = (C.c_double * 2)()
        _native.check(self.lib.pb_solver_xgpu_bench(self._h, iters, mode, edge, ctas, _stream_ptr(None), out),
                      "pb_solver_xgpu_bench")
        return out[0], out[1]

    def tma_probe(self, map_index: int, x: int, y: int, box_cols: int, box_rows: int, local_idx: int = 0,
                  mode: int = 0):
        """Unit-test hook: one TMA box load; returns ``(status, box)`` with ``box`` of shape (box_rows, box_cols)."""
        dev = f"cuda:{self.device}"
        out = torch.full((box_rows, box_cols), float("nan"), dtype=torch.float64, device=dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        _native.check(self.lib.pb_debug_tma_probe(self._h, local_idx, map_index, mode, x, y, box_rows * box_cols * 8,
                                                  out.data_ptr(), status.data_ptr(), _stream_ptr(None)),
                      "pb_debug_tma_probe")
        torch.cuda.synchronize()
        return int(status.item()), out

    # -- kernel-level step API ------------------------------------------------------------------
    def step(self, name: str, *args, local_idx: int = 0, stream=None):
        fn = getattr(self.lib, f"pb_step_{name}")
        if name == "set_options":
            rc = fn(self._h, *args, _stream_ptr(stream))
        else:
            rc = fn(self._h, local_idx, *args, _stream_ptr(stream))
        _native.check(rc, f"pb_step_{name}")

    def read_control(self, local_idx: int = 0) -> dict:
        it, stt, dn = C.c_int(), C.c_int(), C.c_double()
        _native.check(self.lib.pb_step_read_control(self._h, local_idx, C.byref(it), C.byref(stt), C.byref(dn),
                                                    _stream_ptr(None)), "pb_step_read_control")
        return {"iters": it.value, "status": stt.value, "diff_norm": dn.value}


class GroupSolver:
    """Several GPUs driven from this one process (peer access inside the process, one persistent kernel per device):
    the multi-GPU form that needs no launcher.  ``devices`` defaults to ``0..G-1``."""

    def __init__(self, M: int, N: int, Px: int, Py: int, devices=None):
        if not torch.cuda.is_available():
            raise RuntimeError("GroupSolver needs CUDA devices")
        self.lib = _native.lib()
        self.M, self.N, self.Px, self.Py = M, N, Px, Py
        devs = list(range(Px * Py)) if devices is None else list(devices)
        if len(devs) != Px * Py:
            raise ValueError("one device per rank is required")
        torch.cuda.init()
        arr = (C.c_int * len(devs))(*devs)
        self._h = self.lib.pb_group_create(M, N, Px, Py, arr, len(devs))
        if not self._h:
            raise RuntimeError(f"pb_group_create failed: {_native.last_error()}")

    def solve(self, method: str = "persistent", norm: str = "weighted", delta: float = 1e-6, max_iter: int = -1,
              div: str = "markstein", watchdog_s: float = 0.0, tile_rows: int = 0, l2_hints: bool = False,
              full_coefficients: bool = False, ap_mode: int = 0, balance: bool = True, ab_rows: bool = True,
              **_ignored) -> dict:
        opt = _native.SolveOptions(METHODS[method], NORMS[norm], delta, max_iter, DIVS[div], 0, 1, tile_rows, watchdog_s,
                                   int(l2_hints), int(full_coefficients), int(ap_mode), int(not balance), int(not ab_rows))
        st = _native.Stats()
        _native.check(self.lib.pb_group_solve(self._h, C.byref(opt), C.byref(st)), "pb_group_solve")
        out = st.as_dict()
        out["status_name"] = STATUS.get(st.status, "?")
        if st.status in (4, 5):
            raise RuntimeError(f"device-side {STATUS[st.status]}: the solve was abandoned")
        return out

    def solution_host(self) -> torch.Tensor:
        full = torch.zeros((self.M + 1, self.N + 1), dtype=torch.float64)
        _native.check(self.lib.pb_group_copy_solution(self._h, full.data_ptr()), "pb_group_copy_solution")
        return full

    def close(self):
        if getattr(self, "_h", None):
            self.lib.pb_group_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
