#!/usr/bin/env python
"""Per-CTA view of a solve: where every CTA spent its time (phase AB, phase C, waiting in the two reductions) next to
the kinds of tiles it owns.  A CTA that waits long in a reduction finished its phase early; the iteration runs at
the pace of the CTA that never waits.  python bench/experiments/cta_balance.py [M N [max_iter]]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 2 else 4096
    N = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    it = int(sys.argv[3]) if len(sys.argv) > 3 else 400
    method = sys.argv[4] if len(sys.argv) > 4 else "persistent"
    s = GpuSolver(M, N).finalize()
    s.solve(max_iter=it, error_report=False, method=method)
    r = s.solve(max_iter=it, error_report=False, method=method)
    rep = s.cta_report()
    n = r["iters"]
    t = rep[:, :4] / n / 1e3  # us per iteration
    tiles = rep[:, 4:7]
    busy = t[:, 0] + t[:, 1]
    if r["kernel"] == "resident":
        mode, nodes = rep[:, 4], rep[:, 5]
        print(json.dumps({"grid": f"{M}x{N}", "kernel": "resident", "us_per_iter": 1e6 * r["t_iterations"] / n, "ctas": int(rep.shape[0]),
                          "nodes_per_cta_max": int(nodes.max()),
                          "busy_us_uniform_ctas_mean_max": [round(float(busy[mode < 2].mean()), 2), round(float(busy[mode < 2].max()), 2)],
                          "busy_us_mixed_ctas_mean_max": [round(float(busy[mode == 2].mean()), 2), round(float(busy[mode == 2].max()), 2)] if (mode == 2).any() else None,
                          "mixed_ctas": int((mode == 2).sum()),
                          "phase_ab_c_us_cta0": [round(float(t[0, 0]), 2), round(float(t[0, 1]), 2)],
                          "wait_red_us_min": [round(float(t[:, 2].min()), 2), round(float(t[:, 3].min()), 2)]}))
        s.close()
        return
    # least squares: busy time of a CTA = (tiles by kind) . (cost per tile of that kind)
    A = tiles.clone()
    sol = torch.linalg.lstsq(A, busy.unsqueeze(1)).solution.squeeze(1)
    out = {"grid": f"{M}x{N}", "iters": n, "kernel": r["kernel"],
           "fit_us_per_tile": {"recompute_64B": round(float(sol[0]), 3), "stored_80B": round(float(sol[1]), 3), "mixed_112B": round(float(sol[2]), 3)}, "us_per_iter": 1e6 * r["t_iterations"] / n, "ctas": int(rep.shape[0]),
           "tiles_per_cta_min_max": [int(tiles.sum(1).min()), int(tiles.sum(1).max())],
           "mixed_tiles_per_cta_min_mean_max": [int(tiles[:, 2].min()), float(tiles[:, 2].mean()), int(tiles[:, 2].max())]}
    for k, name in enumerate(["phase_ab", "phase_c", "wait_red1", "wait_red2"]):
        out[name + "_us_min_mean_max"] = [round(float(t[:, k].min()), 2), round(float(t[:, k].mean()), 2), round(float(t[:, k].max()), 2)]
    out["busy_us_min_mean_max"] = [round(float(busy.min()), 2), round(float(busy.mean()), 2), round(float(busy.max()), 2)]
    x, y = tiles[:, 2] - tiles[:, 2].mean(), busy - busy.mean()
    out["corr_busy_vs_mixed_tiles"] = float((x * y).sum() / ((x * x).sum() * (y * y).sum()).sqrt().clamp_min(1e-30))
    x = tiles.sum(1) - tiles.sum(1).mean()
    out["corr_busy_vs_tile_count"] = float((x * y).sum() / ((x * x).sum() * (y * y).sum()).sqrt().clamp_min(1e-30))
    order = torch.argsort(busy, descending=True)
    out["slowest"] = [{"cta": int(i), "busy_us": round(float(busy[i]), 2), "ab": round(float(t[i, 0]), 2), "c": round(float(t[i, 1]), 2),
                       "tiles": [int(v) for v in tiles[i]]} for i in order[:6]]
    out["fastest"] = [{"cta": int(i), "busy_us": round(float(busy[i]), 2), "ab": round(float(t[i, 0]), 2), "c": round(float(t[i, 1]), 2),
                       "tiles": [int(v) for v in tiles[i]]} for i in order[-4:]]
    print(json.dumps(out))
    s.close()


if __name__ == "__main__":
    main()
