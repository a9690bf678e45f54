"""Debug aid: exercise single TMA box loads through the solver's tensor maps.
usage: tma_probe.py CASE   (each case in its own process: a faulting load poisons the context)"""
import sys

import torch

sys.path.insert(0, ".")
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402

M, N = 200, 300
s = GpuSolver(M, N).finalize()
s.step("assemble")
torch.cuda.synchronize()
co = s.col_off
cases = {
    "a64_aligned": (4, "a", 64, 33, co + 1, 1),
    "a64_x3": (4, "a", 64, 33, co, 1),
    "a64_x5": (4, "a", 64, 33, co + 2, 1),
    "b66_aligned": (5, "b", 66, 32, co + 1, 1),
    "b66_x3": (5, "b", 66, 32, co, 1),
    "r68x34": (0, "r", 68, 34, co - 1, 0),
    "r68x34_edge": (0, "r", 68, 34, co + 255, 192),
    "a64_neg": (4, "a", 64, 33, -4, -1),
}
name = sys.argv[1]
idx, arrn, bc, br, x, y = cases[name]
arr = s.array(arrn)
st, box = s.tma_probe(idx, x, y, bc, br)
ref = torch.zeros_like(box)
x0, y0 = max(x, 0), max(y, 0)
sub = arr[y0:y + br, x0:x + bc]
ref[y0 - y:y0 - y + sub.shape[0], x0 - x:x0 - x + sub.shape[1]] = sub
print("case", name, "status", st, "match", bool(torch.equal(box, ref)) if st == 1 else None, flush=True)
