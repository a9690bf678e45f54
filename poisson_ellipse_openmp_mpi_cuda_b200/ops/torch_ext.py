"""Loader of the thin PyTorch extension (csrc/binding.cpp): ``load().solve(M, N, ...) -> (report, w)``."""
from __future__ import annotations

import importlib.util
import sys

from .. import _build
from . import _native

_MOD = None


def load():
    """Import the in-tree extension, building it first when it is missing."""
    global _MOD
    if _MOD is not None:
        return _MOD
    import torch  # noqa: F401  (the extension links against libtorch)

    _native.lib()  # libpoisson_b200.so must be loaded (and built) first
    so = _build.PKG / "lib" / "torch_ext" / "poisson_b200_torch.so"
    if not so.exists():
        _MOD = _build.build_torch_extension()
        return _MOD
    spec = importlib.util.spec_from_file_location("poisson_b200_torch", str(so))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sys.modules["poisson_b200_torch"] = mod
    _MOD = mod
    return mod
