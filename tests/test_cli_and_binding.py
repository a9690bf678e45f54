"""CLI golden output, solution dump, PyTorch binding."""
import re
import subprocess

import numpy as np
import pytest
import torch

from poisson_ellipse_openmp_mpi_cuda_b200 import _build
from poisson_ellipse_openmp_mpi_cuda_b200.utils import dump as dump_mod


def _cli(*args):
    _build.build_native()
    return subprocess.run([str(_build.CLI_PATH), *map(str, args)], capture_output=True, text=True, timeout=300)


def test_cli_stage0_format():
    out = _cli("--stage", 0, 40, 40, "--quiet-extras").stdout.splitlines()
    assert out[0] == "Converged after 61 iterations (||w(k+1)-w(k)|| < δ)."
    assert re.fullmatch(r"M=40, N=40 \| Iter=61 \| Time=\d+\.\d{4} s", out[1])


def test_cli_stage0_without_arguments_runs_the_reference_grids():
    out = _cli("--stage", 0, "--quiet-extras").stdout.splitlines()
    assert [ln for ln in out if ln.startswith("Converged")] == [
        f"Converged after {k} iterations (||w(k+1)-w(k)|| < δ)." for k in (17, 31, 61)]
    assert [ln.split(" | Time")[0] for ln in out if ln.startswith("M=")] == [
        "M=10, N=10 | Iter=17", "M=20, N=20 | Iter=31", "M=40, N=40 | Iter=61"]


def test_cli_stage0_grid_list():
    out = _cli("--stage", 0, "--grids", "20x20,10x10", "--quiet-extras").stdout.splitlines()
    assert [ln.split(" | Time")[0] for ln in out if ln.startswith("M=")] == ["M=20, N=20 | Iter=31", "M=10, N=10 | Iter=17"]
    assert _cli("--stage", 0, "--grids", "20x20,10").returncode == 2  # odd number of extents


def test_cli_stage1_format():
    out = _cli("--stage", 1, "--threads", "1,2", 40, 40, "--quiet-extras").stdout.splitlines()
    assert out[:3] == ["--- (Variant 9: Ellipse x^2 + 4y^2 < 1, OpenMP Test) ---", "Grid: M=40, N=40", "-" * 56]
    assert out[3] == out[5] == "Converged after 50 iterations (||w(k+1)-w(k)|| < δ)."
    assert re.fullmatch(r"Threads =  1 \| Time = \d+\.\d{3} s", out[4])
    assert re.fullmatch(r"Threads =  2 \| Time = \d+\.\d{3} s", out[6])
    assert out[7] == "-" * 56


def test_cli_single_positional_argument_is_ignored():
    out = _cli("--stage", 1, 100, "--quiet-extras").stdout
    assert "Converged after 50 iterations" in out  # falls back to 40 x 40 like the reference (argc >= 3)


def test_cli_dump_roundtrip(tmp_path):
    for name in ("w.bin", "w.txt"):
        path = tmp_path / name
        assert _cli("--stage", 1, 20, 20, "--dump", path).returncode == 0
        M, N, arr = dump_mod.read_solution(str(path))
        assert (M, N) == (20, 20) and arr.shape == (21, 21)
        assert arr[0].max() == 0 and arr[:, 0].max() == 0 and arr[-1].max() == 0
        assert arr.max() == pytest.approx(0.09409613374, rel=1e-9)


def test_python_dump_formats(tmp_path):
    import poisson_b200 as pb

    res, w = pb.solve(20, 20, backend="cpu", return_solution=True)
    for name in ("a.bin", "a.npy", "a.dat"):
        p = str(tmp_path / name)
        dump_mod.write_solution(p, w, 20, 20)
        M, N, arr = dump_mod.read_solution(p)
        assert (M, N) == (20, 20)
        assert np.allclose(arr, w.numpy(), rtol=0, atol=0)


@pytest.mark.gpu
def test_cli_stage4_golden():
    out = _cli(800, 1200).stdout.splitlines()
    assert out[0] == "MPI + CUDA 2D run with 1 processes; M=800, N=1200"
    assert out[1] == "Converged after 989 iterations (||w(k+1)-w(k)|| < 1e-06)."
    pats = [r"   GPU compute time \(Ap \+ D\^\{-1\}r, max over ranks\) ~ \S+ s",
            r"   Host<->Device copy time \(max over ranks\)        ~ \S+ s",
            r"   MPI halo exchange time \(max over ranks\)         ~ \S+ s",
            r"   Preconditioner CPU part time \(max over ranks\)   ~ \S+ s",
            r"   Dot products time \(max over ranks\)              ~ \S+ s",
            r"M=800, N=1200 \| Iter=989 \| Total Time=\d+\.\d{6} s",
            r"   Init time \(program\)      ~ \d+\.\d{6} s",
            r"   Solver time \(MPI\+CUDA\)   ~ \d+\.\d{6} s",
            r"   Finalization time        ~ \d+\.\d{6} s"]
    for line, pat in zip(out[2:], pats):
        assert re.fullmatch(pat, line), line
    assert any("L2 error vs analytic (in D)  ~ 1.91571" in ln for ln in out)


@pytest.mark.gpu
def test_cli_virtual_grid_and_dump(tmp_path):
    p = tmp_path / "w.bin"
    r = _cli(400, 600, "--virtual", "--grid", "2x4", "--dump", p)
    assert "MPI + CUDA 2D run with 8 processes; M=400, N=600" in r.stdout
    assert "Iter=546" in r.stdout
    M, N, arr = dump_mod.read_solution(str(p))
    assert arr.max() == pytest.approx(0.09971655665, rel=1e-6)


@pytest.mark.gpu
def test_torch_binding():
    from poisson_ellipse_openmp_mpi_cuda_b200.ops import torch_ext

    ext = torch_ext.load()
    rep, w = ext.solve(160, 160)
    assert rep["iters"] == 183 and rep["status"] == 1
    assert w.is_cuda and w.dtype == torch.float64 and tuple(w.shape) == (161, 161)
    assert float(w.max()) == pytest.approx(0.09943637621, rel=1e-6)
    rep2, w2 = ext.solve(160, 160, 2, 2)
    assert rep2["iters"] == 183
    assert torch.allclose(w, w2, rtol=0, atol=1e-12)


def test_cli_distributed_cpu_stages_fork_their_own_ranks():
    """--stage 2 / 3: the CLI is its own launcher (fork + shared-memory transport, C++ host loop) and prints the
    reference's stage-2 / stage-3 report (S2:476-497, S3:486-507)."""
    import subprocess

    from poisson_ellipse_openmp_mpi_cuda_b200 import _build

    _build.build_native()
    for stage, ranks, extra, banner in [(2, 4, [], "Pure MPI 2D run with 4 processes; M=160, N=160"),
                                        (3, 2, ["--threads", "2"], "MPI/OpenMP run with 2 MPI processes; M=160, N=160"),
                                        (2, 6, ["--grid", "3x2"], "Pure MPI 2D run with 6 processes; M=160, N=160")]:
        out = subprocess.run([str(_build.CLI_PATH), "160", "160", "--stage", str(stage), "--ranks", str(ranks), *extra],
                             capture_output=True, text=True, timeout=120)
        assert out.returncode == 0, out.stderr
        lines = out.stdout.splitlines()
        assert lines[0] == banner
        assert lines[1] == "Converged after 183 iterations (||w(k+1)-w(k)|| < 1e-06)."
        assert lines[2].startswith("M=160, N=160 | Iter=183 | Time=")
        assert "L2 error vs analytic (in D)  ~ 8.9567" in out.stdout
