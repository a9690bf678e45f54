"""Solver families (the reference's stages) and their exact stdout formats."""
import re

from poisson_ellipse_openmp_mpi_cuda_b200.models import FAMILIES, REFERENCE_PROGRAMS, run_family, run_program
from poisson_ellipse_openmp_mpi_cuda_b200.utils import report


def test_stage0_family_text():
    res, text = run_family("stage0", 40, 40)
    lines = text.splitlines()
    assert res["iters"] == 61
    assert lines[0] == "Converged after 61 iterations (||w(k+1)-w(k)|| < δ)."
    assert re.fullmatch(r"M=40, N=40 \| Iter=61 \| Time=\d+\.\d{4} s", lines[1])


def test_stage1_and_stage2_family_text():
    res, text = run_family("stage1", 40, 40, threads=1)
    assert res["iters"] == 50
    assert text.splitlines()[0] == "Converged after 50 iterations (||w(k+1)-w(k)|| < δ)."  # S1:192 prints the symbol
    assert re.fullmatch(r"Threads =  1 \| Time = \d+\.\d{3} s", text.splitlines()[1])
    res, text = run_family("stage2", 20, 20)
    lines = text.splitlines()
    assert lines[0] == "Pure MPI 2D run with 1 processes; M=20, N=20"
    assert lines[1] == "Converged after 26 iterations (||w(k+1)-w(k)|| < 1e-06)."
    assert re.fullmatch(r"M=20, N=20 \| Iter=26 \| Time=\d+\.\d{6} s", lines[2])
    _, text3 = run_family("stage3", 20, 20, threads=2)
    assert text3.splitlines()[0] == "MPI/OpenMP run with 1 MPI processes; M=20, N=20"


def test_stage0_program_sweeps_the_reference_grids():
    assert REFERENCE_PROGRAMS["stage0"]["grids"] == [(10, 10), (20, 20), (40, 40)]  # S0:177
    results, text = run_program("stage0")
    assert [r["iters"] for r in results] == [17, 31, 61]
    lines = text.splitlines()
    assert len(lines) == 6
    for k, (g, it) in enumerate(((10, 17), (20, 31), (40, 61))):
        assert lines[2 * k] == f"Converged after {it} iterations (||w(k+1)-w(k)|| < δ)."
        assert re.fullmatch(rf"M={g}, N={g} \| Iter={it} \| Time=\d+\.\d{{4}} s", lines[2 * k + 1])


def test_stage1_program_prints_banner_sweep_and_rule():
    assert REFERENCE_PROGRAMS["stage1"] == {"grid": (800, 1200), "threads": [4, 8, 16, 32]}  # S1:204, S1:213
    results, text = run_program("stage1-v1", threads=[1, 2])
    assert [r["iters"] for r in results] == [50, 50]
    lines = text.splitlines()
    assert lines[:3] == ["--- (Variant 9: Ellipse x^2 + 4y^2 < 1, OpenMP Test) ---", "Grid: M=40, N=40", "-" * 56]
    assert lines[3] == lines[5] == "Converged after 50 iterations (||w(k+1)-w(k)|| < δ)."
    assert re.fullmatch(r"Threads =  1 \| Time = \d+\.\d{3} s", lines[4])
    assert re.fullmatch(r"Threads =  2 \| Time = \d+\.\d{3} s", lines[6])
    assert lines[7] == "-" * 56 and len(lines) == 8


def test_stage4_report_is_verbatim():
    res = {"iters": 989, "status": 1, "t_gpu": 0.526, "t_copy": 0.0208, "t_comm": 0.00194, "t_precond": 0.1288,
           "t_dot": 0.1668}
    text = report.stage4_report(800, 1200, 2, res, 1e-6, t_init=0.000026, t_solver=0.839, extras=False)
    assert text.splitlines() == [
        "MPI + CUDA 2D run with 2 processes; M=800, N=1200",
        "Converged after 989 iterations (||w(k+1)-w(k)|| < 1e-06).",
        "   GPU compute time (Ap + D^{-1}r, max over ranks) ~ 0.526 s",
        "   Host<->Device copy time (max over ranks)        ~ 0.0208 s",
        "   MPI halo exchange time (max over ranks)         ~ 0.00194 s",
        "   Preconditioner CPU part time (max over ranks)   ~ 0.1288 s",
        "   Dot products time (max over ranks)              ~ 0.1668 s",
        "M=800, N=1200 | Iter=989 | Total Time=0.839026 s",
        "   Init time (program)      ~ 0.000026 s",
        "   Solver time (MPI+CUDA)   ~ 0.839000 s",
        "   Finalization time        ~ 0.000000 s",
    ]


def test_every_family_names_its_reference_stage():
    assert set(FAMILIES) == {"stage0", "stage1", "stage2", "stage3", "stage4", "stage4-nccl"}
    for name, (backend, norm, doc) in FAMILIES.items():
        assert backend in ("cpu", "cpu_dist", "fused", "nccl") and norm in ("weighted", "unweighted") and doc
