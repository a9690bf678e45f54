"""Where recomputing Ap in phase C starts to pay (development aid): stored vs recomputed per block size.

One JSON line per grid: [us/iteration, phase AB, phase C, reduction 1, reduction 2] for ap_mode 1 and 2, best of 2,
200 iterations.  The default rule in Solver::ap_recompute (five arrays > 80 % of L2) comes from this sweep.
"""
import json, sys
import torch
sys.path.insert(0, ".")
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver
for M, N in [(769, 4097), (1025, 4097), (1537, 4097), (2049, 4097), (8192, 8192)]:
    s = GpuSolver(M, N)
    out = {}
    for ap_mode in (1, 2):
        best = None
        for _ in range(2):
            r = s.solve(tile_rows=32, ap_mode=ap_mode, max_iter=200, error_report=False)
            if best is None or r["t_iterations"] < best["t_iterations"]:
                best = r
        it = max(1, best["iters"])
        out[ap_mode] = [round(1e6 * best[k] / it, 2) for k in ("t_iterations", "t_phase_ab", "t_phase_c", "t_red1", "t_red2")]
    print(json.dumps({"grid": [M, N], "nodes_M": round((M - 1) * (N - 1) / 1e6, 2), "stored": out[1], "recomputed": out[2]}), flush=True)
    s.close()
    torch.cuda.empty_cache()
