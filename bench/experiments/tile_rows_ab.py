#!/usr/bin/env python
"""Tile height x Ap mode on the block shapes of multi-GPU ranks, one GPU, same process (source of the auto rules in
Solver::solve_persistent).  python bench/experiments/tile_rows_ab.py [rounds]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402


def main():
    rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    shapes = [(512, 4096, 1500), (1024, 4096, 1500), (2048, 4096, 1000), (800, 1200, 989)]
    variants = [(0, 0), (32, 1), (64, 1), (32, 2), (64, 2)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for M, N, iters in shapes:
        s = GpuSolver(M, N)
        s.solve(method="persistent_tma", max_iter=50, error_report=False)
        for rnd in range(rounds):
            for tile_rows, ap_mode in variants:
                flush.fill_(rnd)
                r = s.solve(method="persistent_tma", max_iter=iters, tile_rows=tile_rows, ap_mode=ap_mode, error_report=False)
                it = r["iters"]
                print(json.dumps({"M": M, "N": N, "tile_rows": tile_rows, "ap_mode": ap_mode, "iters": it,
                                  "us_it": round(1e6 * r["t_iterations"] / it, 2),
                                  "ab": round(1e6 * r["t_phase_ab"] / it, 2), "c": round(1e6 * r["t_phase_c"] / it, 2),
                                  "r1": round(1e6 * r["t_red1"] / it, 2), "r2": round(1e6 * r["t_red2"] / it, 2),
                                  "tiles": s.tile_counts()}), flush=True)
        s.close()
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
