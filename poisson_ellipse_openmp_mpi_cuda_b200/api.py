"""Public Python API: ``solve(M, N, ...)`` and the reusable ``PoissonSolver``.

Backends
--------
``fused``    GPU, the product: persistent cooperative kernel, peer-memory halos, in-kernel reductions.  Uses
             every rank of the current ``torch.distributed`` group (one process per GPU), or one GPU when
             there is no group.
``group``    GPU, same kernels, but all ``gpus`` devices driven from this one process (no launcher needed).
``virtual``  one GPU emulating a ``Px x Py`` process grid inside one cooperative launch (tests, what-if runs).
``nccl``     GPU, library-collective baseline (step kernels + NCCL send/recv + all-reduce).
``cpu``      whole-domain CPU solver (serial or OpenMP): the reference's stages 0/1 and our oracle.
``cpu_dist`` distributed CPU solver over gloo: the reference's stages 2/3.
``auto``     ``fused`` when CUDA is available, else ``cpu``.

The reference has no library API (its only interface is ``argv``: S4:995-999); ``solve`` takes the same
``M, N`` and fixes ``delta = 1e-6`` and ``max_iter = (M-1)(N-1)`` by default like S4:1005-1006.
"""
from __future__ import annotations

import ctypes as C
import time

import torch

from .ops import _native
from .parallel.decomp import parse_grid
from .parallel.fabric import world_info

_NORM = {"weighted": 0, "unweighted": 1}


class PoissonSolver:
    """A solver bound to one grid size and backend; ``solve()`` may be called repeatedly."""

    def __init__(self, M: int, N: int, backend: str = "auto", grid=None, threads: int = 1, gpus: int | None = None):
        if M < 2 or N < 2:
            raise ValueError("M and N must be >= 2")
        if backend == "auto":
            backend = "fused" if torch.cuda.is_available() else "cpu"
        self.M, self.N, self.backend, self.threads = M, N, backend, threads
        self.rank, self.world = world_info()
        self._impl = None
        if backend == "fused":
            if self.world > 1:
                from .parallel.fabric import FusedMultiGpuSolver

                self._impl = FusedMultiGpuSolver(M, N, grid)
                self.grid = (self._impl.Px, self._impl.Py)
            else:
                from .ops.gpu import GpuSolver

                self._impl = GpuSolver(M, N).finalize()
                self.grid = (1, 1)
        elif backend == "group":
            from .ops.gpu import GroupSolver
            from .parallel.decomp import parse_grid as _pg

            n = torch.cuda.device_count() if gpus is None else int(gpus)
            devices = None
            if grid is None or grid == "auto":
                from .parallel.topology import grid_from_topology, probe_p2p

                px, py, devices, _ = grid_from_topology(probe_p2p(range(n))["perf"], M)
                self.grid = (px, py)
            else:
                self.grid = _pg(grid, n, M)
            self._impl = GroupSolver(M, N, self.grid[0], self.grid[1], devices)
        elif backend == "virtual":
            from .ops.gpu import GpuSolver

            if grid is None:
                raise ValueError("backend='virtual' needs grid=(Px, Py)")
            if isinstance(grid, str):
                grid = tuple(int(v) for v in grid.lower().split("x"))
            self.grid = (int(grid[0]), int(grid[1]))
            self._impl = GpuSolver(M, N, self.grid[0], self.grid[1]).finalize()
        elif backend in ("nccl", "cpu_dist"):
            from .parallel.collective import CollectiveSolver

            self._impl = CollectiveSolver(M, N, grid, device="cuda" if backend == "nccl" else "cpu", threads=threads)
            self.grid = (self._impl.Px, self._impl.Py)
        elif backend == "cpu":
            self.grid = (1, 1)
            self._w = None
        else:
            raise ValueError(f"unknown backend {backend!r}")

    # ------------------------------------------------------------------------------------------
    def solve(self, norm: str = "weighted", delta: float = 1e-6, max_iter: int | None = None,
              method: str = "persistent", div: str = "markstein", **kw) -> dict:
        """Run one solve; returns the report (iterations, status, norms, errors, timers)."""
        mi = -1 if max_iter is None else int(max_iter)
        t0 = time.perf_counter()
        if self.backend in ("fused", "virtual", "group"):
            res = self._impl.solve(method=method, norm=norm, delta=delta, max_iter=mi, div=div, **kw)
        elif self.backend in ("nccl", "cpu_dist"):
            res = self._impl.solve(norm=norm, delta=delta, max_iter=mi, div=div)
        else:
            st = _native.Stats()
            self._w = torch.zeros((self.M + 1, self.N + 1), dtype=torch.float64)
            _native.check(_native.lib().pb_cpu_solve(self.M, self.N, _NORM[norm], delta, mi, self.threads,
                                                     int(kw.get("signed_breakdown", 0)), self._w.data_ptr(),
                                                     C.byref(st)), "pb_cpu_solve")
            res = st.as_dict()
            res["status_name"] = {0: "max_iter", 1: "converged", 2: "breakdown"}[st.status]
            res["t_iterations"] = st.seconds
        res["wall_seconds"] = time.perf_counter() - t0
        res.update(M=self.M, N=self.N, backend=self.backend, grid=list(self.grid), nranks=self.world, delta=delta)
        return res

    def solution(self) -> torch.Tensor:
        """Full ``(M+1, N+1)`` nodal field on the host (zero on the boundary of the rectangle).

        Multi-process backends gather the blocks to every rank with ``all_gather_object``; this is a
        convenience for dumps and tests, not a hot path (the reference discards its solution, S4:945-953).
        """
        if self.backend == "cpu":
            if self._w is None:
                raise RuntimeError("solve() first")
            return self._w
        if self.backend in ("virtual", "group") or (self.backend == "fused" and self.world == 1):
            return self._impl.solution_host()
        import torch.distributed as dist

        if self.backend == "fused":
            blk = self._impl.block()
            mine = self._impl.solver.gather_block(0).cpu()
        else:
            blk = self._impl.blk
            mine = self._impl.solution_interior().cpu()
        parts = [None] * self.world
        dist.all_gather_object(parts, (blk.i0, blk.j0, mine))
        full = torch.zeros((self.M + 1, self.N + 1), dtype=torch.float64)
        for i0, j0, t in parts:
            full[i0:i0 + t.shape[0], j0:j0 + t.shape[1]] = t
        return full

    def close(self):
        if self._impl is not None and hasattr(self._impl, "close"):
            self._impl.close()
        self._impl = None


_CACHE: dict = {}


def get_solver(M: int, N: int, backend: str = "auto", grid=None, threads: int = 1, gpus: int | None = None):
    """The cached :class:`PoissonSolver` that :func:`solve` uses for these arguments (created on first use)."""
    if gpus is not None and gpus > 1 and backend in ("auto", "fused") and world_info()[1] == 1:
        backend = "group"  # several GPUs, one process
    key = (M, N, backend, str(grid), threads, gpus, world_info())
    solver = _CACHE.get(key)
    if solver is None:
        solver = _CACHE[key] = PoissonSolver(M, N, backend, grid, threads, gpus)
    return solver


def solve(M: int, N: int, backend: str = "auto", grid=None, norm: str = "weighted", delta: float = 1e-6,
          max_iter: int | None = None, threads: int = 1, dump: str | None = None, return_solution: bool = False,
          gpus: int | None = None, **kw):
    """Solve the fictitious-domain Poisson problem on an ``M x N`` grid.

    Returns the report dict, or ``(report, w)`` with ``return_solution=True``.  Solvers are cached per
    ``(M, N, backend, grid, threads)`` so repeated calls reuse device memory and peer mappings.
    """
    solver = get_solver(M, N, backend, grid, threads, gpus)
    res = solver.solve(norm=norm, delta=delta, max_iter=max_iter, **kw)
    w = None
    if dump or return_solution:
        w = solver.solution()
    if dump and solver.rank == 0:
        from .utils.dump import write_solution

        write_solution(dump, w, M, N)
    return (res, w) if return_solution else res


def clear_cache():
    for s in _CACHE.values():
        s.close()
    _CACHE.clear()
