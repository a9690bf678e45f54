#!/usr/bin/env python
"""Distributed CPU stages against the reference's own programs (C26; published table R4 p.10).

ours      poisson_b200 M N --stage 2 --ranks P             (C++ host loop, shared-memory transport, fork launcher)
          poisson_b200 M N --stage 3 --ranks P --threads T
reference baseline/_ref/bin/stage2_mpi, stage3_hybrid: the unmodified sources (S2, S3) over the MPI shim
Both print their own barrier-to-barrier "Time="; that number is compared.  One JSON line per configuration.

    python bench/cpu_stages.py [--M 800 --N 1200] [--configs 4x1,8x1,4x2]      # ranks x threads
"""
import argparse
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))
import build_ref  # noqa: E402
import run_ref  # noqa: E402


def grab(txt):
    m = re.search(r"Iter=(\d+) \| Time=([0-9.]+) s", txt)
    return (int(m.group(1)), float(m.group(2))) if m else (None, None)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", type=int, default=800)
    ap.add_argument("--N", type=int, default=1200)
    ap.add_argument("--configs", default="4x1,8x1,4x2")
    a = ap.parse_args()
    build_ref.build(["stage2_mpi", "stage3_hybrid"])
    cli = os.path.join(ROOT, "poisson_ellipse_openmp_mpi_cuda_b200", "bin", "poisson_b200")
    for cfg in a.configs.split(","):
        ranks, threads = (int(v) for v in cfg.split("x"))
        stage = 2 if threads == 1 else 3
        ours = subprocess.run([cli, str(a.M), str(a.N), "--stage", str(stage), "--ranks", str(ranks), "--threads", str(threads)],
                              capture_output=True, text=True)
        env_threads = dict(OMP_NUM_THREADS=str(threads))
        os.environ.update(env_threads)
        ref = run_ref.run("stage2_mpi" if stage == 2 else "stage3_hybrid", ranks, [a.M, a.N], bind_gpus=False)
        (it_o, t_o), (it_r, t_r) = grab(ours.stdout), grab(ref)
        print(json.dumps({"grid": f"{a.M}x{a.N}", "stage": stage, "ranks": ranks, "threads": threads, "cpus": os.cpu_count(),
                          "ours_s": t_o, "ours_iters": it_o, "reference_s": t_r, "reference_iters": it_r,
                          "speedup": (t_r / t_o) if t_o and t_r else None}), flush=True)
