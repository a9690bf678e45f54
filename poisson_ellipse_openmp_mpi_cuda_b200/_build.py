"""In-tree native build: nvcc/g++ -> lib/libpoisson_b200.so, bin/poisson_b200, (optional) torch extension.

Replaces the reference's Makefile (stage4-mpi+cuda/Makefile:1-30, which targets sm_60 and names a source
file that is not shipped).  Everything is compiled for sm_100a only:
``-gencode arch=compute_100a,code=sm_100a -lineinfo``.  Objects go to ``build/`` (git-ignored), products
stay inside the package so that they travel with a snapshot of the repository.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = ROOT / "csrc"
BUILD = ROOT / "build"
LIB_DIR = PKG / "lib"
BIN_DIR = PKG / "bin"
LIB_PATH = LIB_DIR / "libpoisson_b200.so"
CLI_PATH = BIN_DIR / "poisson_b200"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
HOST_FLAGS = ["-O3", "-fPIC", "-fopenmp", "-ffp-contract=off", "-Wall", "-Wno-unknown-pragmas"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr", "-rdc=false"]

LIB_SOURCES = [
    "core/decomp.cpp",
    "cpu/oracle.cpp",
    "cpu/dist_solver.cpp",
    "runtime/shm_comm.cpp",
    "api/c_api_cpu.cpp",
    "cuda/assemble.cu",
    "cuda/step_kernels.cu",
    "cuda/persistent.cu",
    "cuda/persistent_tma.cu",
    "cuda/resident.cu",
    "cuda/apply_a_op.cu",
    "runtime/solver.cu",
    "runtime/solver_run.cu",
    "runtime/group.cu",
    "api/c_api_cuda.cu",
]
CLI_SOURCES = ["cli/main.cpp"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _deps_newer(obj: Path, src: Path) -> bool:
    if not obj.exists():
        return True
    t = obj.stat().st_mtime
    if src.stat().st_mtime > t:
        return True
    for hdr in list(CSRC.rglob("*.h")) + list(CSRC.rglob("*.cuh")) + list((ROOT / "include").glob("*.h")):
        if hdr.stat().st_mtime > t:
            return True
    return Path(__file__).stat().st_mtime > t


def _compile(src_rel: str, verbose: bool) -> Path:
    src = CSRC / src_rel
    obj = BUILD / (src_rel.replace("/", "_") + ".o")
    if not _deps_newer(obj, src):
        return obj
    cmd = [_nvcc(), *NVCC_FLAGS, *ARCH_FLAGS, "-Xcompiler", ",".join(HOST_FLAGS), "-I", str(ROOT / "include"),
           "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
    if src.suffix == ".cu" and verbose:
        cmd += ["-Xptxas", "-v"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"compile failed: {' '.join(cmd)}\n{res.stdout}\n{res.stderr}")
    if verbose and res.stderr.strip():
        (BUILD / (obj.name + ".ptxas.txt")).write_text(res.stderr)
    return obj


def existing_sources(names):
    """The listed sources; a name that is not on disk is an error (a library silently built without one of its
    translation units would only fail at load time, on the GPU box)."""
    missing = [s for s in names if not (CSRC / s).exists()]
    if missing:
        raise FileNotFoundError(f"native sources listed in _build.py are missing: {missing}")
    return list(names)


def build_native(verbose: bool = False, force: bool = False) -> Path:
    """Compile the shared library and the CLI; returns the library path."""
    BUILD.mkdir(exist_ok=True)
    LIB_DIR.mkdir(exist_ok=True)
    BIN_DIR.mkdir(exist_ok=True)
    if force:
        for o in BUILD.glob("*.o"):
            o.unlink()
    lib_src = existing_sources(LIB_SOURCES)
    cli_src = existing_sources(CLI_SOURCES)
    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), lib_src + cli_src))
    lib_objs, cli_objs = objs[: len(lib_src)], objs[len(lib_src):]
    need_link = (not LIB_PATH.exists()) or any(o.stat().st_mtime > LIB_PATH.stat().st_mtime for o in lib_objs)
    if need_link:
        cmd = [_nvcc(), "-shared", *ARCH_FLAGS, "-o", str(LIB_PATH), *map(str, lib_objs), "-Xcompiler",
               "-fopenmp", "-lgomp"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed: {' '.join(cmd)}\n{res.stdout}\n{res.stderr}")
    if cli_objs:
        need_cli = (not CLI_PATH.exists()) or need_link or any(
            o.stat().st_mtime > CLI_PATH.stat().st_mtime for o in cli_objs)
        if need_cli:
            cmd = [_nvcc(), *ARCH_FLAGS, "-o", str(CLI_PATH), *map(str, cli_objs), "-L", str(LIB_DIR),
                   "-lpoisson_b200", "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../lib", "-Xcompiler", "-fopenmp",
                   "-lgomp"]
            res = subprocess.run(cmd, capture_output=True, text=True)
            if res.returncode != 0:
                raise RuntimeError(f"CLI link failed: {' '.join(cmd)}\n{res.stdout}\n{res.stderr}")
    return LIB_PATH


def build_torch_extension(verbose: bool = False):
    """Build the thin PyTorch binding (csrc/binding.cpp) in-tree and return the imported module."""
    from torch.utils.cpp_extension import load

    build_native(verbose=verbose)
    ext_dir = PKG / "lib" / "torch_ext"
    ext_dir.mkdir(parents=True, exist_ok=True)
    return load(
        name="poisson_b200_torch",
        sources=[str(CSRC / "binding.cpp")],
        extra_include_paths=[str(ROOT / "include"), "/usr/local/cuda/include"],
        extra_cflags=["-O3", "-std=c++17"],
        extra_ldflags=[f"-L{LIB_DIR}", "-lpoisson_b200", f"-Wl,-rpath,{LIB_DIR}", "-L/usr/local/cuda/lib64",
                       "-lcudart"],
        build_directory=str(ext_dir),
        with_cuda=False,
        verbose=verbose,
    )


if __name__ == "__main__":
    p = build_native(verbose="-v" in sys.argv, force="--force" in sys.argv)
    print(p)
    if "--torch" in sys.argv:
        build_torch_extension(verbose="-v" in sys.argv)
