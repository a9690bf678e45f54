"""The two build descriptions (poisson_ellipse_openmp_mpi_cuda_b200/_build.py, CMakeLists.txt) must name the same
sources, and every one of them must exist: a translation unit missing from one list builds a library that only fails
at load time, on the GPU box."""
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def test_cmake_and_python_build_name_the_same_existing_sources():
    from poisson_ellipse_openmp_mpi_cuda_b200 import _build

    py = {"csrc/" + s for s in _build.LIB_SOURCES}
    cmake = set(re.findall(r"^\s+(csrc/\S+\.(?:cu|cpp))\s*\)?$", (ROOT / "CMakeLists.txt").read_text(), flags=re.M))
    cmake.discard("csrc/cli/main.cpp")
    assert py == cmake, (sorted(py - cmake), sorted(cmake - py))
    for rel in sorted(py):
        assert (ROOT / rel).is_file(), rel
    script = (ROOT / "bench" / "experiments" / "build_variant.sh").read_text()
    for rel in sorted(py):
        assert rel[len("csrc/"):] in script, f"build_variant.sh does not compile {rel}"
