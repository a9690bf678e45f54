"""Debug aid / evidence: last-bit agreement of the cheap division sequences with IEEE division."""
import sys

import torch

sys.path.insert(0, ".")
from poisson_ellipse_openmp_mpi_cuda_b200.ops import _native  # noqa: E402

lib = _native.lib()
torch.cuda.init()
for name, h in [("h1 M=4096", 2.0 / 4096), ("h2 N=4096", 1.2 / 4096), ("h1 M=800", 2.0 / 800), ("h2 N=1200", 1.2 / 1200),
                ("h1 M=333", 2.0 / 333), ("h2 N=517", 1.2 / 517), ("h2 N=600", 1.2 / 600)]:
    c = torch.zeros(4, dtype=torch.int64, device="cuda")
    lib.pb_debug_div_check(h, 200_000_000, 12345, c.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    n, b1, b2, b0 = c.tolist()
    print(f"{name}: samples {n}  x*y mismatches {b0}  one-correction {b1}  two-correction {b2}", flush=True)
