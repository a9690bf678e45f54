"""Dev aid: multi-GPU overhead at fixed local size.  torchrun --nproc-per-node G bench/proxy2.py ROWS_PER_GPU N ITERS"""
import os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import poisson_b200 as pb
from poisson_ellipse_openmp_mpi_cuda_b200.parallel.fabric import init_distributed
rank, world, local = init_distributed("nccl")
rows, N, it = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
s = pb.PoissonSolver(rows * world, N, backend="fused", grid=sys.argv[4] if len(sys.argv) > 4 else "auto")
for rep in range(3):
    r = s.solve(max_iter=it)
    if rank == 0:
        n = r["iters"]
        print(world, s.grid, "us/iter", round(1e6 * r["t_iterations"] / n, 2), "ab", round(1e6 * r["t_phase_ab"] / n, 2), "c",
              round(1e6 * r["t_phase_c"] / n, 2), "red1", round(1e6 * r["t_red1"] / n, 2), "red2", round(1e6 * r["t_red2"] / n, 2), flush=True)
s.close()
if dist.is_initialized():
    dist.barrier(); dist.destroy_process_group()
