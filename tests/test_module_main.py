"""``python -m poisson_b200``: the launcher-side entry point (stages 0-3 on CPU here; stage 4 needs a GPU)."""
import re
import subprocess
import sys

import pytest

from poisson_ellipse_openmp_mpi_cuda_b200.utils import dump as dump_mod


def _run(*args, launcher=()):
    cmd = [sys.executable, *launcher, "-m", "poisson_b200", *map(str, args)]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=300)


def test_stage0_program_without_arguments():
    out = _run("--stage", 0).stdout.splitlines()
    assert [ln for ln in out if ln.startswith("M=")][-1].startswith("M=40, N=40 | Iter=61 | Time=")
    assert out[0] == "Converged after 17 iterations (||w(k+1)-w(k)|| < δ)."


def test_stage1_thread_sweep():
    out = _run("--stage", 1, 40, 40, "--threads", "1,2").stdout.splitlines()
    assert out[0].startswith("--- (Variant 9") and out[1] == "Grid: M=40, N=40" and out[-1] == "-" * 56
    assert sum(ln.startswith("Threads = ") for ln in out) == 2


def test_stage2_single_process_and_single_positional():
    out = _run(20, 20, "--stage", 2, "--quiet-extras").stdout.splitlines()
    assert out[0] == "Pure MPI 2D run with 1 processes; M=20, N=20"
    assert out[1] == "Converged after 26 iterations (||w(k+1)-w(k)|| < 1e-06)."
    assert re.fullmatch(r"M=20, N=20 \| Iter=26 \| Time=\d+\.\d{6} s", out[2])
    # one positional number is ignored, like the reference: 40 x 40
    assert "M=40, N=40 | Iter=50" in _run(100, "--stage", 2, "--quiet-extras").stdout


def test_stage3_under_torchrun_with_dump(tmp_path):
    path = tmp_path / "w.bin"
    launcher = ("-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                "--master-port", "29593")
    r = _run(40, 40, "--stage", 3, "--threads", 2, "--dump", path, "--quiet-extras", launcher=launcher)
    assert r.returncode == 0, r.stderr[-500:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith(("MPI/OpenMP", "Converged", "M="))]
    assert lines[0] == "MPI/OpenMP run with 2 MPI processes; M=40, N=40"
    assert lines[1] == "Converged after 50 iterations (||w(k+1)-w(k)|| < 1e-06)."
    M, N, arr = dump_mod.read_solution(str(path))
    assert (M, N) == (40, 40) and arr.max() == pytest.approx(0.09797040154, rel=1e-9)


@pytest.mark.gpu
def test_stage4_single_gpu():
    out = _run(400, 600).stdout.splitlines()
    assert out[0] == "MPI + CUDA 2D run with 1 processes; M=400, N=600"
    assert out[1] == "Converged after 546 iterations (||w(k+1)-w(k)|| < 1e-06)."
    assert any(ln.startswith("M=400, N=600 | Iter=546 | Total Time=") for ln in out)
