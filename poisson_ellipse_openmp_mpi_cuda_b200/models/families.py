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


def run_family(name: str, M: int, N: int, threads: int = 1, grid=None, delta: float = 1e-6, **kw):
    backend, norm, _ = FAMILIES[name]
    rank, world = world_info()
    t_prog = time.perf_counter()
    solver = api.PoissonSolver(M, N, backend=backend, grid=grid, threads=threads)
    t_before = time.perf_counter()
    extra = {"signed_breakdown": 1} if name in ("stage0", "stage1") else {}
    res = solver.solve(norm=norm, delta=delta, **extra, **kw)
    t_after = time.perf_counter()
    if name == "stage0":
        text = report.stage0_report(M, N, res, res["seconds"])
    elif name == "stage1":
        text = report.stage1_report(M, N, threads, res, delta, res["seconds"])
    elif name in ("stage2", "stage3"):
        text = report.stage23_report(M, N, world, res, delta, res["seconds"], hybrid=(name == "stage3"))
    else:
        text = report.stage4_report(M, N, world, res, delta, t_init=t_before - t_prog, t_solver=t_after - t_before)
    solver.close()
    return res, text
