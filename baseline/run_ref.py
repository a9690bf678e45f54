"""Launch a reference program on P shim ranks (one process per rank, CUDA_VISIBLE_DEVICES = rank for stage 4).

    python baseline/run_ref.py --prog poisson_mpi_cuda --ranks 2 800 1200
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
import time
from pathlib import Path

BIN = Path(__file__).resolve().parent / "_ref" / "bin"


def run(prog: str, ranks: int, args, bind_gpus: bool = True, timeout: float = 3600.0) -> str:
    """Returns rank 0's stdout."""
    job = f"{os.getpid()}_{time.time_ns()}"
    procs = []
    for r in range(ranks):
        env = dict(os.environ, PB_MPI_RANK=str(r), PB_MPI_SIZE=str(ranks), PB_MPI_JOB=job)
        if bind_gpus:
            env["CUDA_VISIBLE_DEVICES"] = str(r)
        procs.append(subprocess.Popen([str(BIN / prog), *map(str, args)], env=env, text=True,
                                      stdout=subprocess.PIPE if r == 0 else subprocess.DEVNULL))
    out, _ = procs[0].communicate(timeout=timeout)
    for p in procs[1:]:
        p.wait(timeout=timeout)
    if any(p.returncode for p in procs):
        raise RuntimeError(f"{prog} failed: {[p.returncode for p in procs]}\n{out}")
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--prog", default="poisson_mpi_cuda")
    ap.add_argument("--ranks", type=int, default=1)
    ap.add_argument("--no-gpu-bind", action="store_true")
    ap.add_argument("args", nargs="*")
    a = ap.parse_args()
    sys.stdout.write(run(a.prog, a.ranks, a.args, bind_gpus=not a.no_gpu_bind))
