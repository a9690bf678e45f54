#!/usr/bin/env python
"""Soak of the fused multi-GPU path: N solves in a row on all ranks, every one must converge in the same number of
iterations to the same bits (stop norm, error norms).  A slip in the halo-record / proxy-fence / reduction protocol
shows up as a different iteration count, different bits or a watchdog status.

    python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1 bench/tools/soak.py [--solves 200] [--M 4096 --N 4096]
"""
import argparse
import json
import os
import sys
import time

import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import poisson_b200 as pb  # noqa: E402
from poisson_ellipse_openmp_mpi_cuda_b200.parallel.fabric import init_distributed  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--solves", type=int, default=200)
    ap.add_argument("--M", type=int, default=4096)
    ap.add_argument("--N", type=int, default=4096)
    a = ap.parse_args()
    rank, world, _ = init_distributed("nccl")
    s = pb.PoissonSolver(a.M, a.N, backend="fused")
    first, bad, t0, dev = None, [], time.perf_counter(), 0.0
    for i in range(a.solves):
        r = s.solve(watchdog_s=10.0)
        key = (r["iters"], r["status_name"], r["diff_norm"].hex(), float(r["l2_in_D"]).hex(), float(r["max_abs"]).hex())
        dev += r["seconds"]
        if first is None:
            first = key
        elif key != first:
            bad.append((i, key))
    out = {"what": "soak", "n_gpus": world, "grid": f"{a.M}x{a.N}", "process_grid": list(s.grid), "solves": a.solves,
           "iterations": first[0], "status": first[1], "all_identical": not bad, "differing_solves": bad[:5],
           "device_seconds_per_solve": dev / a.solves, "wall_seconds": time.perf_counter() - t0}
    if rank == 0:
        print(json.dumps(out), flush=True)
    s.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0 if not bad else 1


if __name__ == "__main__":
    sys.exit(main())
