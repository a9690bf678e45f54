"""Tiny single solve (debug aid): python bench/tiny_solve.py M N method tile_rows Px Py"""
import sys

sys.path.insert(0, ".")
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402

M, N = int(sys.argv[1]), int(sys.argv[2])
method = sys.argv[3] if len(sys.argv) > 3 else "persistent"
tile_rows = int(sys.argv[4]) if len(sys.argv) > 4 else 0
Px = int(sys.argv[5]) if len(sys.argv) > 5 else 1
Py = int(sys.argv[6]) if len(sys.argv) > 6 else 1
s = GpuSolver(M, N, Px, Py)
r = s.solve(method=method, tile_rows=tile_rows, max_iter=int(sys.argv[7]) if len(sys.argv) > 7 else -1)
print({k: r[k] for k in ("iters", "status_name", "seconds", "l2_in_D", "ctas_per_rank")})
