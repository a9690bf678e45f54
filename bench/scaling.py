#!/usr/bin/env python
"""Secondary benchmark configurations of BASELINE.json (one JSON line each, rank 0):

  * 800 x 1200 to convergence (989 iterations)                     -- config 2
  * 4096 x 4096 to convergence (3226 iterations)                   -- config 3 (same as bench.py)
  * weak scaling, 8192 x 8192 nodes per GPU, fixed 200 iterations  -- config 4 (us/iteration and fraction of the
    HBM roofline: 112 B/node/iteration at the measured copy bandwidth)

    python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1 bench/scaling.py [--grid reference]

Device-timed (CUDA events inside the solver), max over ranks, best of `--reps` after one warm-up solve.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import poisson_b200 as pb  # noqa: E402
from poisson_ellipse_openmp_mpi_cuda_b200.parallel.fabric import init_distributed  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", default="auto")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--skip-weak", action="store_true")
    a = ap.parse_args()
    rank, world, local = init_distributed("nccl")
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0)) * 1e9
    configs = [("800x1200", 800, 1200, None), ("4096x4096", 4096, 4096, None)]
    if not a.skip_weak:
        configs.append((f"weak_8192x8192_per_gpu", 8192 * world, 8192, 200))
    for name, M, N, max_iter in configs:
        s = pb.PoissonSolver(M, N, backend="fused", grid=a.grid)
        best = None
        for rep in range(a.reps + 1):
            r = s.solve(max_iter=max_iter)
            if rep and (best is None or r["seconds"] < best["seconds"]):
                best = r
        us = 1e6 * best["t_iterations"] / best["iters"]
        nodes = (M - 1) * (N - 1)
        out = {"config": name, "M": M, "N": N, "n_gpus": world, "grid": f"{s.grid[0]}x{s.grid[1]}", "iters": best["iters"],
               "status": best["status_name"], "seconds": best["seconds"], "us_per_iter": us,
               "l2_error_in_D": best["l2_in_D"], "roofline_us_per_iter": 1e6 * 112.0 * nodes / world / hbm,
               "fraction_of_hbm_roofline": (112.0 * nodes / world / hbm) / (us * 1e-6),
               "fraction_of_hbm_roofline_80B_moved": (80.0 * nodes / world / hbm) / (us * 1e-6),
               "timers_s": {k: best[k] for k in ("t_assemble", "t_gpu", "t_comm", "t_dot")}}
        if rank == 0:
            print(json.dumps(out), flush=True)
        s.close()
        torch.cuda.empty_cache()
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
