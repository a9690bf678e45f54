"""Target of bench/tools/sanitize.sh: the whole multi-rank protocol (halo records, proxy fences, record-star
reductions) on one GPU as 2 x 2 virtual ranks, 160 x 160, plus a single-rank solve with several tiles per CTA."""
import sys

sys.path.insert(0, ".")
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402

s = GpuSolver(160, 160, 2, 2)
r = s.solve(watchdog_s=600.0)
print("virtual 2x2 160x160:", r["iters"], r["status_name"], r["l2_in_D"])
assert r["iters"] == 183
s.close()
s = GpuSolver(400, 600)
r = s.solve(max_iter=20, watchdog_s=600.0)
print("single 400x600, 20 iterations:", r["iters"], r["status_name"])
s.close()
