#!/usr/bin/env python
"""Stages 0 (serial) and 1 (OpenMP) against the reference's own programs on the CPUs of this machine.

reference  baseline/_ref/bin/stage0_v2 (400 x 600), stage1_v2 (400 x 600; sweeps 2/4/8/16 threads itself),
           stage1_v3 (800 x 1200; 4/8/16/32 threads): the unmodified sources, grids and thread counts hard-coded in them
ours       poisson_b200 M N --stage 0 | --stage 1 --threads T
Both print "Time=" measured around the solver call; that number is compared.  One JSON line per configuration.

    python bench/cpu_stage01.py [--big]      # --big adds 800 x 1200
"""
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref", "bin")
CLI = os.path.join(ROOT, "poisson_ellipse_openmp_mpi_cuda_b200", "bin", "poisson_b200")


def run(cmd):
    return subprocess.run(cmd, capture_output=True, text=True).stdout


def serial_time(txt):
    m = re.search(r"Iter=(\d+) \| Time=([0-9.]+) s", txt)
    return int(m.group(1)), float(m.group(2))


def thread_times(txt):
    iters = [int(v) for v in re.findall(r"Converged after (\d+) iterations", txt)]
    return {int(t): float(s) for t, s in re.findall(r"Threads =\s*(\d+) \| Time = ([0-9.]+) s", txt)}, (iters[0] if iters else None)


def ours_threads(M, N, t, reps=2):
    best = None
    for _ in range(reps):
        times, iters = thread_times(run([CLI, str(M), str(N), "--stage", "1", "--threads", str(t)]))
        best = times[t] if best is None else min(best, times[t])
    return best, iters


def main():
    out = []
    it_r, t_r = serial_time(run([os.path.join(REF, "stage0_v2")]))
    it_o, t_o = min((serial_time(run([CLI, "400", "600", "--stage", "0"])) for _ in range(2)), key=lambda v: v[1])
    out.append({"stage": 0, "grid": "400x600", "reference_s": t_r, "reference_iters": it_r, "ours_s": t_o, "ours_iters": it_o})
    configs = [("stage1_v2", 400, 600)] + ([("stage1_v3", 800, 1200)] if "--big" in sys.argv else [])
    for prog, M, N in configs:
        ref, it_r = thread_times(run([os.path.join(REF, prog)]))
        for t, t_r in sorted(ref.items()):
            if t > (os.cpu_count() or 1):
                continue  # more threads than CPUs: the reference's own sweep goes there, the comparison does not
            t_o, it_o = ours_threads(M, N, t)
            out.append({"stage": 1, "grid": f"{M}x{N}", "threads": t, "reference_s": t_r, "reference_iters": it_r,
                        "ours_s": t_o, "ours_iters": it_o})
    for o in out:
        o["cpus"] = os.cpu_count()
        o["speedup"] = round(o["reference_s"] / o["ours_s"], 2)
        print(json.dumps(o), flush=True)


if __name__ == "__main__":
    main()
