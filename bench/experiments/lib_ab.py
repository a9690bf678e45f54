#!/usr/bin/env python
"""A/B of native-library variants (build/variants/NAME, see build_variant.sh; "." = the in-tree library) on the block
shapes of 1-, 2-, 4- and 8-GPU ranks of 4096 x 4096, one GPU, interleaved processes in one gpurun call.
python bench/experiments/lib_ab.py "NAME NAME ..." [rounds]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def one(tag):
    import torch
    from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for M, N, iters in [(4096, 4096, 3226), (2048, 4096, 1500), (1024, 4096, 1500), (512, 4096, 1500), (800, 1200, 989)]:
        s = GpuSolver(M, N)
        s.solve(method="persistent_tma", max_iter=50, error_report=False)
        best = None
        for rep in range(2):
            flush.fill_(rep)
            r = s.solve(method="persistent_tma", max_iter=iters, error_report=False)
            if best is None or r["t_iterations"] < best["t_iterations"]:
                best = r
        it = best["iters"]
        print(json.dumps({"lib": tag, "M": M, "N": N, "iters": it, "status": best["status_name"],
                          "us_it": round(1e6 * best["t_iterations"] / it, 2), "ab": round(1e6 * best["t_phase_ab"] / it, 2),
                          "c": round(1e6 * best["t_phase_c"] / it, 2), "r1": round(1e6 * best["t_red1"] / it, 2),
                          "r2": round(1e6 * best["t_red2"] / it, 2), "diff": best["diff_norm"]}), flush=True)
        s.close()
        torch.cuda.empty_cache()


def main():
    if sys.argv[1] == "--one":
        one(sys.argv[2])
        return
    libs = sys.argv[1].split()
    rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    for _ in range(rounds):
        for name in libs:
            env = dict(os.environ)
            if name == ".":
                env.pop("POISSON_B200_LIB", None)
            else:
                env["POISSON_B200_LIB"] = os.path.join(ROOT, "build", "variants", name, "libpoisson_b200.so")
            p = subprocess.run([sys.executable, os.path.abspath(__file__), "--one", name], env=env, capture_output=True, text=True, timeout=400)
            sys.stdout.write(p.stdout)
            if p.returncode != 0:
                print(json.dumps({"lib": name, "failed": p.stderr[-400:]}))
            sys.stdout.flush()


if __name__ == "__main__":
    main()
