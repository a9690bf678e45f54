"""Build the reference's own programs, unmodified, into baseline/_ref/ (git-ignored, shipped to the GPU box).

`pip install /root/reference` is impossible (the reference is five stand-alone C++/CUDA programs with no
setup.py / pyproject.toml -- see DESIGN.md), so "installing the reference" means: copy its sources verbatim
to baseline/_ref/src/ and compile them with the flags of its Makefile (stage4-mpi+cuda/Makefile:6-13:
``nvcc -O3 -std=c++11 -arch=...``), retargeted to sm_100, against the MPI shim in baseline/mpishim/ because
the image has no MPI.  The three files that do not compile as shipped get the one-token ``-D`` fix
described in SURVEY.md section 0.4 on the command line; no source file is edited.
"""
from __future__ import annotations

import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = Path("/root/reference")
DST = HERE / "_ref"
SRC = DST / "src"
BIN = DST / "bin"
SHIM = HERE / "mpishim"

PROGRAMS = {
    # name: (source relative to the reference root, kind, extra flags)
    "stage0_v1": ("stage0/Withoutopenmp1.cpp", "cpp", []),
    "stage0_v2": ("stage0/Withoutopenmp2.cpp", "cpp", ["-Depsilon=eps"]),
    "stage0_v3": ("stage0/Withoutopenmp3.cpp", "cpp", ["-Depsilon=eps"]),
    "stage1_v1": ("stage1-openmp/Withopenmp1.cpp", "omp", ["-Dep=eps"]),
    "stage1_v2": ("stage1-openmp/Withopenmp2.cpp", "omp", []),
    "stage1_v3": ("stage1-openmp/Withopenmp3.cpp", "omp", []),
    "stage2_mpi": ("stage2-mpi/poisson_mpi_decomp.cpp", "mpi", []),
    "stage3_hybrid": ("stage3-mpi+openmp/main_hybrid.cpp", "mpi_omp", []),
    "poisson_mpi_cuda": ("stage4-mpi+cuda/poisson_mpi_cuda2.cu", "cuda", []),
    # micro-benchmark driver (baseline/ref_comm_bench.cu) around the unmodified stage-4 source, which it #includes
    "ref_comm_bench": ("stage4-mpi+cuda/poisson_mpi_cuda2.cu", "cuda_bench", []),
}


def sync_sources() -> bool:
    """Copy the reference sources (only possible where /root/reference exists)."""
    if not REF.exists():
        return SRC.exists()
    for _, (rel, _, _) in PROGRAMS.items():
        out = SRC / rel
        out.parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(REF / rel, out)
    return True


def build(names=None, verbose=False) -> dict:
    BIN.mkdir(parents=True, exist_ok=True)
    if not sync_sources():
        raise RuntimeError("reference sources unavailable (no /root/reference and no baseline/_ref/src)")
    results = {}
    for name, (rel, kind, extra) in PROGRAMS.items():
        if names and name not in names:
            continue
        src, out = SRC / rel, BIN / name
        if out.exists() and out.stat().st_mtime > src.stat().st_mtime and out.stat().st_mtime > (
                SHIM / "mpishim.cpp").stat().st_mtime:
            results[name] = "up to date"
            continue
        if kind == "cuda_bench":
            cmd = ["nvcc", "-O3", "-std=c++11", "-arch=sm_100", "-DPB_SHIM_WITH_CUDA", f"-I{SHIM}", f"-I{SRC}",
                   str(HERE / "ref_comm_bench.cu"), str(SHIM / "mpishim.cpp"), "-o", str(out), "-lrt"]
        elif kind == "cuda":
            cmd = ["nvcc", "-O3", "-std=c++11", "-arch=sm_100", "-DPB_SHIM_WITH_CUDA", f"-I{SHIM}", str(src),
                   str(SHIM / "mpishim.cpp"), "-o", str(out), "-lrt"]
        else:
            cmd = ["g++", "-O3", "-std=c++11", *extra, str(src), "-o", str(out)]
            if kind in ("omp", "mpi_omp"):
                cmd.append("-fopenmp")
            if kind in ("mpi", "mpi_omp"):
                cmd += [f"-I{SHIM}", str(SHIM / "mpishim.cpp"), "-lrt", "-pthread"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        results[name] = "ok" if res.returncode == 0 else f"FAILED: {res.stderr[-400:]}"
        if verbose:
            print(name, results[name])
    return results


if __name__ == "__main__":
    r = build(sys.argv[1:] or None, verbose=True)
    sys.exit(0 if all(v in ("ok", "up to date") for v in r.values()) else 1)
