#!/usr/bin/env python
"""Multi-GPU overhead at a fixed per-GPU block: split of the iteration into phases and reduction waits.

    python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1 bench/proxy2.py ROWS_PER_GPU N ITERS [GRID]

Solves a (ROWS_PER_GPU * G) x N grid for ITERS iterations on G GPUs (or one GPU without torchrun), so that every GPU
count works on the same local block; prints us/iteration of phase AB, phase C and the two reductions (device clocks,
max over ranks).  Used to see what the neighbours and the cross-GPU reductions add to a single-GPU iteration.
"""
import os
import sys

import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import poisson_b200 as pb  # noqa: E402
from poisson_ellipse_openmp_mpi_cuda_b200.parallel.fabric import init_distributed  # noqa: E402


def main():
    rank, world, _ = init_distributed("nccl")
    rows, N, iters = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    s = pb.PoissonSolver(rows * world, N, backend="fused", grid=sys.argv[4] if len(sys.argv) > 4 else "auto")
    for _ in range(3):
        r = s.solve(max_iter=iters)
        if rank == 0:
            n = r["iters"]
            us = {k: round(1e6 * r[k] / n, 2) for k in ("t_iterations", "t_phase_ab", "t_phase_c", "t_red1", "t_red2")}
            print(world, s.grid, us, flush=True)
    s.close()
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
