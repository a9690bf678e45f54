"""Per-CTA phase clocks of one persistent solve (development aid): how evenly the CTAs finish their phases.

Usage: python bench/cta_times.py M N [max_iter]  -> min / median / max over the CTAs of the per-iteration time (us)
spent in phase AB, phase C and waiting in the two reductions, plus the ten CTAs that wait least in reduction 1
(= the ones the others wait for).
"""
import json
import sys

import torch

sys.path.insert(0, ".")
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402

M, N = int(sys.argv[1]), int(sys.argv[2])
max_iter = int(sys.argv[3]) if len(sys.argv) > 3 else 150
s = GpuSolver(M, N)
s.solve(max_iter=max_iter, error_report=False)
r = s.solve(max_iter=max_iter, error_report=False)
t = s.cta_times() * 1e6 / max(1, r["iters"])
q = lambda v: [round(float(x), 2) for x in (v.min(), v.median(), v.max())]
out = {"grid": [M, N], "iters": r["iters"], "ctas": int(t.shape[0]), "ab": q(t[:, 0]), "c": q(t[:, 1]),
       "wait_red1": q(t[:, 2]), "wait_red2": q(t[:, 3]),
       "slowest_in_ab": [int(i) for i in torch.argsort(t[:, 2])[:10]],
       "slowest_in_c": [int(i) for i in torch.argsort(t[:, 3])[:10]]}
print(json.dumps(out), flush=True)
s.close()
