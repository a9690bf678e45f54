"""The ctypes mirrors in ops/_native.py must match the structs of include/poisson_b200.h field by field.

A struct that grows on one side only is read past its end on the other (this is a C ABI, nothing checks it at run
time), so the header is parsed here and compared with the Python declarations, and every exported function the
Python layer declares must exist in the built library.
"""
import ctypes as C
import re
from pathlib import Path

import pytest

from poisson_ellipse_openmp_mpi_cuda_b200.ops import _native

HEADER = Path(__file__).resolve().parents[1] / "include" / "poisson_b200.h"
CTYPES = {"int": C.c_int, "double": C.c_double, "long long": C.c_longlong}


def _struct_fields(name: str):
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), text, flags=re.S).group(1)
    fields = []
    for decl in body.split(";"):
        decl = " ".join(decl.split())
        if not decl:
            continue
        m = re.fullmatch(r"(long long|int|double) (.+)", decl)
        assert m, decl
        for item in m.group(2).split(","):
            item = item.strip()
            arr = re.fullmatch(r"(\w+)\[(\d+)\]", item)
            if arr:
                fields.append((arr.group(1), CTYPES[m.group(1)] * int(arr.group(2))))
            else:
                fields.append((item, CTYPES[m.group(1)]))
    return fields


@pytest.mark.parametrize("c_name,mirror", [("pb_stats", _native.Stats), ("pb_block", _native.Block),
                                           ("pb_solve_options", _native.SolveOptions)])
def test_struct_mirrors_match_the_header(c_name, mirror):
    want = _struct_fields(c_name)
    got = list(mirror._fields_)
    assert [n for n, _ in got] == [n for n, _ in want]
    for (name, t_got), (_, t_want) in zip(got, want):
        assert C.sizeof(t_got) == C.sizeof(t_want) and t_got._type_ == t_want._type_, name


def test_every_declared_function_is_exported_and_declared_in_the_header():
    lib = _native.lib()
    text = HEADER.read_text()
    declared = set(re.findall(r"\b(pb_\w+)\s*\(", text))
    # the functions the Python layer binds (restype / argtypes set by _native._declare)
    bound = [n for n in dir(lib) if n.startswith("pb_")] or sorted(declared)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} is in the header but not exported by the library"
    for name in bound:
        assert name in declared, f"{name} is bound in Python but missing from the header"
