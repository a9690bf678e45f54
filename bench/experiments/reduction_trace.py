#!/usr/bin/env python
"""Where the in-kernel reduction spends its time: clock64 stamps between the steps of star_sum (micro-benchmark mode
5 of pb_solver_xgpu_bench), for CTA 0 (which gathers) and for the other CTAs.  One GPU:
python bench/experiments/reduction_trace.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver  # noqa: E402

STEPS = ["shuffles+sts+proxy_fence", "barrier", "warp_sum", "release_fence+store", "gather(cta0)",
         "xrank+fence+result_store(cta0)", "result_poll", "acquire_load"]


def main():
    s = GpuSolver(2048, 2048).finalize()
    iters = 2000
    for mode, name in [(0, "counter+aggregator"), (3, "record star")]:
        s.xgpu_bench(50, mode, 0)
        ns, chk = s.xgpu_bench(iters, mode, 0)
        print(json.dumps({"mode": mode, "what": name, "us_per_reduction": ns / 1e3, "checksum_ok": chk == iters * 148}))
    s.xgpu_bench(50, 5, 0)
    ns, _ = s.xgpu_bench(iters, 5, 0)
    rep = s.cta_report()
    print(json.dumps({"mode": 5, "what": "record star, traced", "us_per_reduction": ns / 1e3,
                      "cycles_cta0": dict(zip(STEPS, [round(float(v)) for v in rep[0]])),
                      "cycles_others_mean": dict(zip(STEPS, [round(float(v)) for v in rep[1:].mean(dim=0)])),
                      "cycles_others_max": dict(zip(STEPS, [round(float(v)) for v in rep[1:].max(dim=0).values]))}))
    s.close()


if __name__ == "__main__":
    main()
