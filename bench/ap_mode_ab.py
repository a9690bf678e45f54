"""A/B of the two ways phase C gets Ap on constant-coefficient tiles (development aid).

ap_mode 1: phase AB stores Ap, phase C reloads it (80 B/node/iteration).  ap_mode 2: phase C evaluates the
stencil again on the p_new box (about 66 B/node/iteration, one more stencil evaluation per node).
Usage: python bench/ap_mode_ab.py [max_iter]   -> one JSON line per (grid, tile_rows, ap_mode), best of 2.
"""
import json
import sys

import torch

sys.path.insert(0, ".")
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402


def main():
    max_iter = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    for M, N in [(4096, 4096), (513, 4097), (800, 1200)]:
        s = GpuSolver(M, N)
        for tile_rows in (32, 16):
            for ap_mode in (1, 2):
                best = None
                for _ in range(2):
                    r = s.solve(tile_rows=tile_rows, ap_mode=ap_mode, max_iter=max_iter, error_report=False)
                    if best is None or r["t_iterations"] < best["t_iterations"]:
                        best = r
                it = max(1, best["iters"])
                print(json.dumps({"grid": [M, N], "tile_rows": tile_rows, "ap_mode": ap_mode, "iters": best["iters"],
                                  "us_per_iter": round(1e6 * best["t_iterations"] / it, 2),
                                  "ab": round(1e6 * best["t_phase_ab"] / it, 2), "c": round(1e6 * best["t_phase_c"] / it, 2),
                                  "red1": round(1e6 * best["t_red1"] / it, 2), "red2": round(1e6 * best["t_red2"] / it, 2),
                                  "diff_norm": best["diff_norm"]}), flush=True)
        s.close()
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
