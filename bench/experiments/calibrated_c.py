#!/usr/bin/env python
"""A/B of the calibrated phase-C assignment against the static one, one GPU, same process, on the block shapes the
ranks of a 1-, 2- and 4-GPU run of 4096 x 4096 work on.  python bench/experiments/calibrated_c.py [rounds]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402


def main():
    rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    shapes = [(4096, 4096, 3226), (2048, 4096, 1500), (1024, 4096, 1500)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for M, N, iters in shapes:
        s = GpuSolver(M, N)
        s.solve(method="persistent_tma", max_iter=100, error_report=False)  # measures the pace
        pace = s.phase_c_pace()
        print(json.dumps({"M": M, "N": N, "pace_ns_per_unit_min_mean_max": [round(float(pace.min()), 2), round(float(pace.mean()), 2),
                                                                            round(float(pace.max()), 2)] if pace.numel() else None}), flush=True)
        for rnd in range(rounds):
            for balance in ("static", True):
                flush.fill_(rnd)
                r = s.solve(method="persistent_tma", max_iter=iters, balance=balance, error_report=False)
                it = r["iters"]
                rep = s.cta_report()
                tc = rep[:, 1] / it / 1e3
                print(json.dumps({"M": M, "N": N, "balance": str(balance), "iters": it, "status": r["status_name"],
                                  "us_it": round(1e6 * r["t_iterations"] / it, 2),
                                  "ab": round(1e6 * r["t_phase_ab"] / it, 2), "c": round(1e6 * r["t_phase_c"] / it, 2),
                                  "r1": round(1e6 * r["t_red1"] / it, 2), "r2": round(1e6 * r["t_red2"] / it, 2),
                                  "c_us_min_mean_max_over_ctas": [round(float(tc.min()), 2), round(float(tc.mean()), 2), round(float(tc.max()), 2)],
                                  "diff": r["diff_norm"]}), flush=True)
        s.close()
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
