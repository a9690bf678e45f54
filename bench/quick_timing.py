"""Quick device-timed sweep of the single-GPU solver (development aid; not the driver's bench)."""
import json
import sys

import torch

sys.path.insert(0, ".")
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402


def main():
    grids = [(800, 1200), (4096, 4096)] if len(sys.argv) < 3 else [(int(sys.argv[1]), int(sys.argv[2]))]
    for M, N in grids:
        s = GpuSolver(M, N)
        for method in ("persistent",):
            for exact in (("markstein",) if len(sys.argv) > 3 else ("true", "markstein", "recip")):
                for cps in ((32, 16) if method == "persistent" else (2,) if method == "persistent_ldg" else (0,)):
                    best = None
                    for rep in range(2):
                        r = s.solve(method=method, div=exact, ctas_per_sm=cps if method != "persistent" else 0,
                                    tile_rows=cps if method == "persistent" else 0)
                        if best is None or r["seconds"] < best["seconds"]:
                            best = r
                    print(json.dumps({"grid": [M, N], "method": method, "exact": exact, "ctas_per_sm_or_tile_rows": cps,
                                      "ctas": best["ctas_per_rank"], "iters": best["iters"],
                                      "seconds": round(best["seconds"], 5),
                                      "us_per_iter": round(1e6 * best["t_iterations"] / max(1, best["iters"]), 2),
                                      "t_asm_ms": round(1e3 * best["t_assemble"], 3),
                                      "t_gpu": round(best["t_gpu"], 4), "t_dot": round(best["t_dot"], 4),
                                      "l2": best["l2_in_D"]}), flush=True)
        s.close()
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
