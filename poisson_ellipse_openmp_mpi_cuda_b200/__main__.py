"""``python -m poisson_b200 [M N] [--stage ...]`` -- the launcher-side counterpart of the reference's programs.

The reference is started as ``mpirun -np P ./poisson_mpi_cuda M N`` (S4:985-1038), ``mpirun -np P ./poisson_mpi M N``
(S2:463-502, S3:473-512) or as an argument-less executable (S0:176-196, S1:204-229).  Here the same stages run as

    python -m poisson_b200 800 1200                                   # stage 4 on one GPU
    torchrun --nproc-per-node 8 -m poisson_b200 4096 4096             # stage 4, one process per GPU, fused path
    torchrun --nproc-per-node 8 -m poisson_b200 4096 4096 --stage 4-nccl   # library-collective baseline
    torchrun --nproc-per-node 4 -m poisson_b200 400 600 --stage 3 --threads 2   # MPI+OpenMP analogue (gloo)
    python -m poisson_b200 --stage 0                                  # sequential program: 10^2, 20^2, 40^2
    python -m poisson_b200 --stage 1 --threads 4,8,16,32              # OpenMP program (800 x 1200 unless M N given)

Rank 0 prints the stage's report in the reference's exact format; fewer than two positional numbers mean 40 40
for the distributed stages, as in the reference (argc >= 3, S4:995-999).
"""
from __future__ import annotations

import argparse
import sys


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="python -m poisson_b200", description=__doc__.split("\n\n")[0])
    ap.add_argument("extent", nargs="*", type=int, help="M N (a single number is ignored, like the reference)")
    ap.add_argument("--stage", default="4", choices=["0", "1", "2", "3", "4", "4-nccl"])
    ap.add_argument("--threads", default="1", help="OpenMP threads; stage 1 takes a list: 4,8,16,32")
    ap.add_argument("--grid", default=None, help="process grid PxxPy | reference | auto (stage 4 default: auto)")
    ap.add_argument("--delta", type=float, default=1e-6)
    ap.add_argument("--max-iter", type=int, default=None)
    ap.add_argument("--dump", default=None, help="write the (M+1) x (N+1) solution (.bin, .npy or text)")
    ap.add_argument("--quiet-extras", action="store_true", help="only the reference's own lines")
    a = ap.parse_args(argv)

    from .models import families
    from .parallel.fabric import init_distributed, world_info

    have_mn = len(a.extent) >= 2
    M, N = (a.extent[0], a.extent[1]) if have_mn else (40, 40)
    threads = [int(t) for t in a.threads.split(",") if t]
    kw = {} if a.max_iter is None else {"max_iter": a.max_iter}

    if a.stage == "0":
        _, text = families.run_program("stage0", grids=[(M, N)] if have_mn else None, **kw)
        print(text)
        return 0
    if a.stage == "1":
        _, text = families.run_program("stage1", grid=(M, N) if have_mn else None,
                                       threads=threads if a.threads != "1" or have_mn else None, **kw)
        print(text)
        return 0

    name = "stage" + a.stage
    init_distributed("gloo" if a.stage in ("2", "3") else None)
    rank, world = world_info()
    res, text = families.run_family(name, M, N, threads=threads[0], grid=a.grid, delta=a.delta,
                                    extras=not a.quiet_extras, dump=a.dump, **kw)
    if rank == 0:
        print(text)
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()
    return 3 if res.get("status") == 4 else 0


if __name__ == "__main__":
    sys.exit(main())
