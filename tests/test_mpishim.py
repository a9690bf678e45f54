"""The MPI shim used to run the reference's own MPI programs (baseline/mpishim): real message passing between
processes.  Runs the reference's unmodified stage-2 program on 1, 2 and 4 shim ranks and checks its own report."""
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "baseline"))


@pytest.fixture(scope="module")
def ref_bin():
    import build_ref

    try:
        res = build_ref.build(["stage2_mpi", "stage0_v1"])
    except RuntimeError as e:  # no reference sources on this machine
        pytest.skip(str(e))
    if any(v not in ("ok", "up to date") for v in res.values()):
        pytest.skip(f"reference build failed: {res}")
    return build_ref.BIN


@pytest.mark.parametrize("ranks", [1, 2, 4])
def test_reference_stage2_on_shim_ranks(ref_bin, ranks):
    import run_ref

    out = run_ref.run("stage2_mpi", ranks, [40, 40], bind_gpus=False, timeout=120)
    assert f"Pure MPI 2D run with {ranks} processes; M=40, N=40" in out
    assert "Converged after 50 iterations" in out
    assert "Iter=50" in out


def test_reference_stage0_prints_61(ref_bin):
    import subprocess

    out = subprocess.run([str(ref_bin / "stage0_v1")], capture_output=True, text=True, timeout=60).stdout
    assert "M=40, N=40 | Iter=61" in out and "M=10, N=10 | Iter=17" in out
