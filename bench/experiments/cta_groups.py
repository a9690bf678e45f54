import json, os, sys
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
import torch
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver
tag = sys.argv[1]
for M, N, iters in [(4096, 4096, 1500), (1024, 4096, 1500)]:
    s = GpuSolver(M, N)
    s.solve(method="persistent_tma", max_iter=50, error_report=False)
    r = s.solve(method="persistent_tma", max_iter=iters, error_report=False)
    rep = s.cta_report(); it = r["iters"]
    ab, c = rep[:, 0] / it / 1e3, rep[:, 1] / it / 1e3
    mixed = rep[:, 6]
    out = {"lib": tag, "M": M, "N": N, "us_it": round(1e6 * r["t_iterations"] / it, 2)}
    for m in sorted(set(int(v) for v in mixed.tolist())):
        sel = mixed == m
        out[f"mixed={m}"] = {"ctas": int(sel.sum()), "ab_mean": round(float(ab[sel].mean()), 2), "ab_max": round(float(ab[sel].max()), 2),
                             "c_mean": round(float(c[sel].mean()), 2), "c_max": round(float(c[sel].max()), 2)}
    print(json.dumps(out), flush=True)
    s.close()
