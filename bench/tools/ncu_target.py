"""Target of bench/tools/ncu_capture.sh: a short solve so that one replayed launch of the persistent kernel stays small.
    python bench/tools/ncu_target.py M N iterations [Px Py]"""
import sys

sys.path.insert(0, ".")
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402

M, N, it = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
pg = (int(sys.argv[4]), int(sys.argv[5])) if len(sys.argv) > 5 else (1, 1)
s = GpuSolver(M, N, *pg)
r = s.solve(max_iter=it, error_report=False)
print(r["iters"], r["status_name"], 1e6 * r["t_iterations"] / r["iters"], "us/iteration")
s.close()
