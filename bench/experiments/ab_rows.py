#!/usr/bin/env python
"""A/B of phase AB by warp-private rows (PB_AB_ROWS, read by the library at every launch) against the tile path with
its barrier between the p update and the stencil, with and without the calibrated phase-C assignment; one GPU, same
process, block shapes of the ranks of a 1-, 2-, 4- and 8-GPU run of 4096 x 4096.
python bench/experiments/ab_rows.py [rounds]"""
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
    variants = [("0", "static"), ("1", "static"), ("1", True)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for M, N, iters in shapes:
        s = GpuSolver(M, N)
        os.environ["PB_AB_ROWS"] = "1"
        s.solve(method="persistent_tma", max_iter=100, error_report=False)  # measures the pace of phase C
        for rnd in range(rounds):
            for rows, balance in variants:
                os.environ["PB_AB_ROWS"] = rows
                flush.fill_(rnd)
                r = s.solve(method="persistent_tma", max_iter=iters, balance=balance, error_report=False)
                it = r["iters"]
                rep = s.cta_report()
                ta, tc = rep[:, 0] / it / 1e3, rep[:, 1] / it / 1e3
                print(json.dumps({"M": M, "N": N, "ab_rows": rows, "balance": str(balance), "iters": it, "status": r["status_name"],
                                  "us_it": round(1e6 * r["t_iterations"] / it, 2),
                                  "ab": round(1e6 * r["t_phase_ab"] / it, 2), "c": round(1e6 * r["t_phase_c"] / it, 2),
                                  "r1": round(1e6 * r["t_red1"] / it, 2), "r2": round(1e6 * r["t_red2"] / it, 2),
                                  "ab_us_min_mean_max_over_ctas": [round(float(ta.min()), 2), round(float(ta.mean()), 2), round(float(ta.max()), 2)],
                                  "c_us_min_mean_max_over_ctas": [round(float(tc.min()), 2), round(float(tc.mean()), 2), round(float(tc.max()), 2)],
                                  "diff": r["diff_norm"]}), flush=True)
        s.close()
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
