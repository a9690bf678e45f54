#!/usr/bin/env python
"""Secondary benchmark configurations of BASELINE.json (one JSON line each, rank 0):

  * 800 x 1200 to convergence (989 iterations)                     -- config 2
  * 4096 x 4096 to convergence (3226 iterations)                   -- config 3 (same as bench.py)
  * weak scaling, 8192 x 8192 nodes per GPU, fixed 200 iterations  -- config 4: on the reference's 2 x 4 grid
    (16384 x 32768 nodes) and on 8 x 1 slabs (65536 x 8192 nodes).  The reference itself cannot run this
    configuration: its only options are M and N, max_iter is fixed at (M-1)(N-1) (S4:1006) and the solve would take
    > 10^4 iterations; so this row has no reference column.

    python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1 bench/scaling.py [--configs 800,4096,weak]

Device-timed (CUDA events inside the solver), max over ranks, best of `--reps` after one warm-up solve.  Every line
carries bench.py's roofline block: bytes counted from the kernel's own tile classes, HBM time at the measured copy
bandwidth and at 8 TB/s, halo time over NVLink, reduction latency floor, fraction = model / measured.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as headline  # noqa: E402  (bench.py at the repository root: roofline())
import poisson_b200 as pb  # noqa: E402
from poisson_ellipse_openmp_mpi_cuda_b200.parallel.decomp import choose_process_grid  # noqa: E402
from poisson_ellipse_openmp_mpi_cuda_b200.parallel.fabric import init_distributed  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="800,4096,weak")
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    rank, world, local = init_distributed("nccl")
    dev = torch.device("cuda", local)
    configs = []
    for name in a.configs.split(","):
        if name == "800":
            configs.append(("800x1200", 800, 1200, None, "auto"))
        elif name == "4096":
            configs.append(("4096x4096", 4096, 4096, None, "auto"))
        elif name == "weak":
            px, py = choose_process_grid(world)  # the reference's grid: 2 x 4 at 8 ranks (S4:82-86)
            configs.append((f"weak_8192x8192_per_gpu_{px}x{py}", 8192 * px, 8192 * py, 200, f"{px}x{py}"))
            if world > 1:
                configs.append((f"weak_8192x8192_per_gpu_{world}x1", 8192 * world, 8192, 200, f"{world}x1"))
    for name, M, N, max_iter, grid in configs:
        s = pb.PoissonSolver(M, N, backend="fused", grid=grid)
        best = None
        for rep in range(a.reps + 1):
            r = s.solve(max_iter=max_iter)
            if rep and (best is None or r["seconds"] < best["seconds"]):
                best = r
        gs = s._impl.solver if world > 1 else s._impl
        tc = gs.tile_counts()
        tct = torch.tensor([tc["recompute_64B"], tc["stored_80B"], tc["mixed_112B"]], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tct)
        tiles = dict(zip(("recompute_64B", "stored_80B", "mixed_112B"), (int(v) for v in tct.tolist())))
        us = 1e6 * best["t_iterations"] / best["iters"]
        out = {"config": name, "M": M, "N": N, "n_gpus": world, "grid": f"{s.grid[0]}x{s.grid[1]}", "iters": best["iters"],
               "status": best["status_name"], "seconds": best["seconds"], "us_per_iter": us,
               "l2_error_in_D": best["l2_in_D"], "roofline": headline.roofline(M, N, world, s.grid, tiles, us),
               "timers_s": {k: best[k] for k in ("t_assemble", "t_gpu", "t_comm", "t_dot", "t_phase_ab", "t_phase_c",
                                                  "t_red1", "t_red2")}}
        if rank == 0:
            print(json.dumps(out), flush=True)
        s.close()
        torch.cuda.empty_cache()
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
