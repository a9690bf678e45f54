#!/usr/bin/env python
"""A/B of the phase-C sweep direction (PB_REVERSE_C, read by the library at every launch) and the L2 hint sets on one
GPU, on the block shapes the ranks of a 1-, 2-, 4- and 8-GPU run of 4096 x 4096 work on.

    python bench/experiments/reverse_c.py [rounds]
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402


def main():
    rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    shapes = [(4096, 4096, 3226), (2048, 4096, 1500), (1024, 4096, 1500), (512, 4096, 1500)]
    variants = [("0", 0), ("1", 0), ("1", 2), ("1", 1), ("0", 1)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for M, N, iters in shapes:
        s = GpuSolver(M, N)
        s.solve(method="persistent_tma", max_iter=50, error_report=False)
        for rnd in range(rounds):
            for rev, hints in variants:
                os.environ["PB_REVERSE_C"] = rev
                flush.fill_(rnd)
                r = s.solve(method="persistent_tma", max_iter=iters, l2_hints=hints, error_report=False)
                it = r["iters"]
                print(json.dumps({"M": M, "N": N, "rev": rev, "hints": hints, "iters": it, "status": r["status_name"],
                                  "us_it": round(1e6 * r["t_iterations"] / it, 2),
                                  "ab": round(1e6 * r["t_phase_ab"] / it, 2), "c": round(1e6 * r["t_phase_c"] / it, 2),
                                  "r1": round(1e6 * r["t_red1"] / it, 2), "r2": round(1e6 * r["t_red2"] / it, 2),
                                  "diff": r["diff_norm"]}), flush=True)
        s.close()
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
