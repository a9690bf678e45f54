#!/usr/bin/env python
"""Library-collective baseline (B1 of SURVEY.md section 7.2): the same PCG with NCCL between separate kernel launches.

    python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1 bench/nccl_baseline.py [M N]

Step kernels (csrc/cuda/step_kernels.cu) + `isend/irecv` halo exchange of p + `all_reduce` of the scalars, driven from
Python (about ten collective / kernel calls per iteration, so the loop is host-bound for small blocks -- that is what a
library-collective structure costs).  Prints one JSON line per backend for comparison with the fused path.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import poisson_b200 as pb  # noqa: E402
from poisson_ellipse_openmp_mpi_cuda_b200.parallel.fabric import init_distributed  # noqa: E402


def main():
    rank, world, local = init_distributed("nccl")
    M, N = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (800, 1200)
    for backend, grid in (("nccl", "reference"), ("fused", "reference"), ("fused", "auto")):
        s = pb.PoissonSolver(M, N, backend=backend, grid=grid)
        best = None
        for rep in range(3):
            r = s.solve()
            if rep and (best is None or r["seconds"] < best["seconds"]):
                best = r
        if rank == 0:
            print(json.dumps({"M": M, "N": N, "n_gpus": world, "backend": backend, "grid": f"{s.grid[0]}x{s.grid[1]}",
                              "iters": best["iters"], "seconds": best["seconds"],
                              "us_per_iter": 1e6 * best["t_iterations"] / best["iters"], "l2_error_in_D": best["l2_in_D"]}),
                  flush=True)
        s.close()
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
