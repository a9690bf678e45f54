#!/usr/bin/env python
"""The reference's own halo-exchange and dot + all-reduce paths, micro-benchmarked through the MPI shim
(baseline/ref_comm_bench.cu: the unmodified stage-4 source with its main renamed).  Companion of
bench/halo_allreduce_sweep.py (ours and NCCL); together they are BASELINE.json's config 5.

    python bench/ref_comm_sweep.py --ranks 8 [--iters 200]     # one GPU per rank
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline"))
import build_ref  # noqa: E402
import run_ref  # noqa: E402

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", type=int, default=1)
    ap.add_argument("--iters", type=int, default=200)
    a = ap.parse_args()
    if not (build_ref.BIN / "ref_comm_bench").exists():
        build_ref.build(["ref_comm_bench"])
    sys.stdout.write(run_ref.run("ref_comm_bench", a.ranks, [a.iters]))
