"""Solver families = the reference's five stages, each mapped onto a backend of this framework.

``run_family(name, M, N)`` runs one and returns ``(report_dict, text)`` where ``text`` is the stdout report
in that stage's exact format (utils/report.py).
"""
from __future__ import annotations

import time

from .. import api
from ..parallel.fabric import world_info
from ..utils import report

# name -> (backend, norm, description)
FAMILIES = {
    "stage0": ("cpu", "unweighted", "sequential PCG (S0: Withoutopenmp*.cpp), unweighted stop norm"),
    "stage1": ("cpu", "weighted", "OpenMP PCG (S1: Withopenmp*.cpp)"),
    "stage2": ("cpu_dist", "weighted", "distributed CPU PCG over gloo (S2: poisson_mpi_decomp.cpp)"),
    "stage3": ("cpu_dist", "weighted", "distributed CPU PCG + OpenMP threads (S3: main_hybrid.cpp)"),
    "stage4": ("fused", "weighted", "multi-GPU PCG, fused halo + reductions (S4: poisson_mpi_cuda2.cu)"),
    "stage4-nccl": ("nccl", "weighted", "multi-GPU PCG, NCCL halo + all-reduce (library baseline)"),
}


# what each argument-less reference program runs (S0:177, S0v2:188, S0v3:186; S1v1:204-215, S1v2:204-214, S1:204-214)
REFERENCE_PROGRAMS = {
    "stage0": {"grids": [(10, 10), (20, 20), (40, 40)]},
    "stage0-v2": {"grids": [(400, 600)]},
    "stage0-v3": {"grids": [(800, 1200)]},
    "stage1-v1": {"grid": (40, 40), "threads": [1, 4, 16]},
    "stage1-v2": {"grid": (400, 600), "threads": [2, 4, 8, 16]},
    "stage1": {"grid": (800, 1200), "threads": [4, 8, 16, 32]},
}


def run_program(name: str, grids=None, grid=None, threads=None, **kw):
    """Run a CPU reference program the way its argument-less ``main`` does and return ``(results, stdout text)``.

    Stage 0 sweeps a list of grids (S0:176-196); stage 1 sweeps thread counts on one grid between a banner and a
    closing rule (S1:204-229).  ``grids`` / ``grid`` / ``threads`` override the hard-coded lists.
    """
    spec = REFERENCE_PROGRAMS[name]
    if name.startswith("stage0"):
        out, texts = [], []
        for M, N in (grids or spec["grids"]):
            res, text = run_family("stage0", M, N, extras=False, **kw)
            out.append(res)
            texts.append(text)
        return out, "\n".join(texts)
    M, N = grid or spec["grid"]
    entries = []
    for t in (threads or spec["threads"]):
        res, _ = run_family("stage1", M, N, threads=t, **kw)
        entries.append((t, res, res["seconds"]))
    return [e[1] for e in entries], report.stage1_program(M, N, entries)


def run_family(name: str, M: int, N: int, threads: int = 1, grid=None, delta: float = 1e-6, extras: bool = True,
               dump: str | None = None, **kw):
    """Run one stage's equivalent; returns ``(report dict, stdout text in that stage's format)``.

    ``dump`` writes the full nodal field (collective for the distributed stages; rank 0 writes the file).
    """
    backend, norm, _ = FAMILIES[name]
    rank, world = world_info()
    t_prog = time.perf_counter()
    solver = api.PoissonSolver(M, N, backend=backend, grid=grid, threads=threads)
    t_before = time.perf_counter()
    extra = {"signed_breakdown": 1} if name in ("stage0", "stage1") else {}
    res = solver.solve(norm=norm, delta=delta, **extra, **kw)
    t_after = time.perf_counter()
    if name == "stage0":
        text = report.stage0_report(M, N, res, res["seconds"], extras=extras)
    elif name == "stage1":
        text = report.stage1_report(M, N, threads, res, delta, res["seconds"])
    elif name in ("stage2", "stage3"):
        text = report.stage23_report(M, N, world, res, delta, res["seconds"], hybrid=(name == "stage3"), extras=extras)
    else:
        text = report.stage4_report(M, N, world, res, delta, t_init=t_before - t_prog, t_solver=t_after - t_before,
                                    extras=extras)
    if dump:
        from ..utils.dump import write_solution

        w = solver.solution()
        if rank == 0:
            write_solution(dump, w, M, N)
    solver.close()
    return res, text
