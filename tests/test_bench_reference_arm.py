"""bench.py's reference arm, exercised without a GPU: a canned stage-4 report (the format of S4:900-902,
S4:969-978, S4:1026-1033) goes through the same parser and line builder the real arm uses, and the line must
carry every key the driver reads.  (Round 1 lost its only judged number to a NameError in this function.)"""
import json
import math
import os
import stat
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import bench  # noqa: E402

CANNED = """MPI + CUDA 2D run with {P} processes; M=4096, N=4096
Converged after 3226 iterations (||w(k+1)-w(k)|| < 1e-06).
   GPU compute time (Ap + D^{{-1}}r, max over ranks) ~ 2.85653 s
   Host<->Device copy time (max over ranks)        ~ 0.0636763 s
   MPI halo exchange time (max over ranks)         ~ 0.000428541 s
   Preconditioner CPU part time (max over ranks)   ~ 0.574033 s
   Dot products time (max over ranks)              ~ 1.17067 s
M=4096, N=4096 | Iter=3226 | Total Time={total:.6f} s
   Init time (program)      ~ 0.000082 s
   Solver time (MPI+CUDA)   ~ {solver:.6f} s
   Finalization time        ~ 0.000000 s
"""

REQUIRED = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
            "vs_baseline", "dtype", "data", "config", "impl", "samples", "iterations", "us_per_iter", "timers_s",
            "e2e", "clocks", "gpu_launches")


def test_parser_reads_every_number_of_the_report():
    r = bench.parse_reference_report(CANNED.format(P=1, total=4.7626, solver=4.762518))
    assert r["iters"] == 3226 and r["converged"]
    assert r["solver_s"] == 4.762518 and r["total_s"] == 4.7626 and r["init_s"] == 0.000082
    assert r["t_gpu"] == 2.85653 and r["t_copy"] == 0.0636763 and r["t_comm"] == 0.000428541
    assert r["t_precond"] == 0.574033 and r["t_dot"] == 1.17067


def test_reference_line_has_every_key_and_an_e2e_block():
    reps = [bench.parse_reference_report(CANNED.format(P=2, total=3.2 + 0.1 * i, solver=3.1 + 0.1 * i)) for i in range(3)]
    line = bench.reference_line(reps, [5.0, 5.1, 5.2], 4096, 4096, 2, steps=3, warmup=3, clocks=None)
    json.dumps(line)  # serialisable
    for k in REQUIRED:
        assert k in line, k
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["unit"] == "s"
    assert math.isclose(line["value"], 3.2) and line["samples"] == [3.1, 3.2, 3.3]
    assert line["iterations"] == 3226 and line["expected_iterations"] == 3226
    e = line["e2e"]
    assert e["unit"] == line["unit"] and math.isclose(e["value"], 3.3) and e["value"] != line["value"]
    assert math.isclose(e["process_wall_s"], 5.1)
    assert e["h2d_bytes_per_step"] > 3 * 2048 * 4096 * 8 and e["d2h_bytes_per_step"] > 3226 * 3 * 262144
    # same keys as our own arm prints for the ratio the driver computes
    for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert k in e


def test_copy_byte_model_follows_the_reference_source():
    h2d, d2h = bench.reference_copy_bytes(800, 1200, 1, 989)
    assert h2d == 3 * 801 * 1201 * 8  # three (nx+2)(ny+2) arrays, no halo copies on one rank
    assert d2h == 262144 * (1 + 3 * 989)


def test_reference_arm_end_to_end_with_a_stand_in_program(tmp_path):
    """The whole `--impl reference` path (process launch, parsing, one JSON line on stdout) against a stand-in
    executable that prints a canned report, so that no part of the arm is first executed on the GPU box."""
    fake_root = tmp_path / "repo"
    (fake_root / "baseline" / "_ref" / "bin").mkdir(parents=True)
    exe = fake_root / "baseline" / "_ref" / "bin" / "poisson_mpi_cuda"
    exe.write_text("#!/bin/sh\ncat <<'EOF'\n" + CANNED.format(P=1, total=4.7626, solver=4.762518) + "EOF\n")
    exe.chmod(exe.stat().st_mode | stat.S_IEXEC)
    (fake_root / "bench.py").write_text((ROOT / "bench.py").read_text())
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    p = subprocess.run([sys.executable, str(fake_root / "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1"],
                       capture_output=True, text=True, env=env, timeout=120)
    assert p.returncode == 0, p.stderr[-400:]
    lines = [ln for ln in p.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    line = json.loads(lines[0])
    assert "unavailable" not in line, line
    for k in REQUIRED:
        assert k in line, k
    assert line["value"] == 4.762518 and line["e2e"]["value"] == 4.7626 and len(line["samples"]) == 2


def test_roofline_block_counts_bytes_from_tile_classes():
    tiles = {"recompute_64B": 90, "stored_80B": 0, "mixed_112B": 10}
    r = bench.roofline(4096, 4096, 1, (1, 1), tiles, us_iter=230.0)
    assert math.isclose(r["bytes_per_node_iter"], 0.9 * 64 + 0.1 * 112)
    assert r["nvlink_us"] == 0.0 and not r["l2_resident"]
    assert math.isclose(r["model_us"], r["hbm_us"] + r["reduction_floor_us"])
    assert 0.0 < r["fraction_of_roofline"] < 1.5 and r["fraction_of_hbm_bandwidth"] < r["fraction_of_roofline"]
    r8 = bench.roofline(4096, 4096, 8, (8, 1), {"recompute_64B": 0, "stored_80B": 100, "mixed_112B": 0}, us_iter=36.0)
    assert r8["l2_resident"] and r8["nvlink_us"] > 0 and math.isclose(r8["bytes_per_node_iter"], 80.0)
    assert math.isclose(r8["bytes_per_iter_per_gpu"], 80.0 * 4095 * 4095 / 8)
