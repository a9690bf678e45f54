"""ctypes view of libpoisson_b200 (the C ABI in include/poisson_b200.h).

The library is built in-tree by ``_build.build_native`` and loaded from ``<package>/lib``.  There is no
Python fallback for any native function: if the library is missing it is built, and if that fails the
import raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

from .. import _build

_LIB = None


class Stats(C.Structure):
    """Mirror of ``pb_stats``."""

    _fields_ = [
        ("iters", C.c_int),
        ("status", C.c_int),
        ("diff_norm", C.c_double),
        ("l2_in_D", C.c_double),
        ("l2_all", C.c_double),
        ("max_abs", C.c_double),
        ("max_w", C.c_double),
        ("nodes_in_D", C.c_longlong),
        ("seconds", C.c_double),
        ("t_gpu", C.c_double),
        ("t_copy", C.c_double),
        ("t_comm", C.c_double),
        ("t_precond", C.c_double),
        ("t_dot", C.c_double),
        ("t_assemble", C.c_double),
        ("t_iterations", C.c_double),
        ("sum_e2_in_D", C.c_double),
        ("sum_e2_all", C.c_double),
        ("t_phase_ab", C.c_double),
        ("t_phase_c", C.c_double),
        ("t_red1", C.c_double),
        ("t_red2", C.c_double),
    ]

    def as_dict(self) -> dict:
        return {name: getattr(self, name) for name, _ in self._fields_}


class Block(C.Structure):
    """Mirror of ``pb_block``."""

    _fields_ = [
        ("rank", C.c_int),
        ("px", C.c_int),
        ("py", C.c_int),
        ("i0", C.c_int),
        ("j0", C.c_int),
        ("nx", C.c_int),
        ("ny", C.c_int),
        ("pitch", C.c_int),
        ("nbr", C.c_int * 4),
    ]


class SolveOptions(C.Structure):
    """Mirror of ``pb_solve_options``."""

    _fields_ = [
        ("method", C.c_int),
        ("norm", C.c_int),
        ("delta", C.c_double),
        ("max_iter", C.c_longlong),
        ("exact_div", C.c_int),
        ("ctas_per_sm", C.c_int),
        ("error_report", C.c_int),
        ("tile_rows", C.c_int),
        ("watchdog_s", C.c_double),
        ("l2_hints", C.c_int),
        ("full_coefficients", C.c_int),
        ("ap_mode", C.c_int),
        ("no_balance", C.c_int),
        ("ab_mode", C.c_int),
    ]


def _declare(lib: C.CDLL) -> None:
    dp = C.POINTER(C.c_double)
    vp = C.c_void_p
    sig = {
        "pb_version": (C.c_int, []),
        "pb_last_error": (C.c_char_p, []),
        "pb_choose_proc_grid": (None, [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
        "pb_choose_grid_topology": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_int),
                                              C.POINTER(C.c_int), C.POINTER(C.c_int)]),
        "pb_local_block": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(Block)]),
        "pb_col_offset": (C.c_int, []),
        "pb_coef_a": (C.c_double, [C.c_int] * 4),
        "pb_coef_b": (C.c_double, [C.c_int] * 4),
        "pb_rhs": (C.c_double, [C.c_int] * 4),
        "pb_diag": (C.c_double, [C.c_int] * 4),
        "pb_chord": (C.c_double, [C.c_double, C.c_double, C.c_double, C.c_int]),
        "pb_exact": (C.c_double, [C.c_double, C.c_double]),
        "pb_cpu_solve": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_double, C.c_longlong, C.c_int, C.c_int, vp,
                                   C.POINTER(Stats)]),
        "pb_cpu_error_report": (C.c_int, [C.c_int, C.c_int, vp, C.POINTER(Stats)]),
        "pb_cpu_dist_solve": (C.c_int, [C.c_int] * 5 + [C.c_char_p, C.c_int, C.c_double, C.c_longlong, C.c_int, vp,
                                                         C.POINTER(Stats)]),
        "pb_cpu_assemble_local": (C.c_int, [C.c_int, C.c_int, C.POINTER(Block), vp, vp, vp, vp, C.c_int]),
        "pb_cpu_apply_A_local": (C.c_double, [C.c_int, C.c_int, C.POINTER(Block), vp, vp, vp, vp, C.c_int]),
        "pb_cpu_update_wr_local": (C.c_int, [C.POINTER(Block), vp, vp, vp, vp, vp, C.c_double, dp, C.c_int]),
        "pb_cpu_update_p_local": (C.c_int, [C.POINTER(Block), vp, vp, vp, C.c_double, C.c_int]),
        "pb_cpu_dot_zr_local": (C.c_double, [C.POINTER(Block), vp, vp, C.c_int]),
        "pb_cpu_error_local": (C.c_int, [C.c_int, C.c_int, C.POINTER(Block), vp, dp]),
    }
    I, LL, D = C.c_int, C.c_longlong, C.c_double
    sig.update({
        "pb_cuda_device_count": (I, []),
        "pb_runtime_bootstrap": (I, [C.POINTER(I), I]),
        "pb_p2p_matrix": (I, [C.POINTER(I), I, C.POINTER(I), C.POINTER(I), C.POINTER(I)]),
        "pb_solver_create": (vp, [I] * 7),
        "pb_solver_destroy": (None, [vp]),
        "pb_solver_ipc_handle": (I, [vp, I, vp]),
        "pb_solver_open_peer": (I, [vp, I, vp]),
        "pb_solver_set_peer_base": (I, [vp, I, vp]),
        "pb_solver_arena": (vp, [vp, I]),
        "pb_solver_finalize": (I, [vp]),
        "pb_solver_solve": (I, [vp, C.POINTER(SolveOptions), vp, C.POINTER(Stats)]),
        "pb_solver_block": (I, [vp, I, C.POINTER(Block)]),
        "pb_solver_array": (vp, [vp, I, I]),
        "pb_solver_control_red": (vp, [vp, I]),
        "pb_solver_control_err": (vp, [vp, I]),
        "pb_solver_copy_solution": (I, [vp, vp, vp]),
        "pb_solver_gather_block": (I, [vp, I, vp, LL, vp]),
        "pb_solver_info": (I, [vp, C.POINTER(I), C.POINTER(LL), C.POINTER(LL), C.POINTER(LL)]),
        "pb_solver_ap_recompute": (I, [vp, I]),
        "pb_solver_cta_times": (I, [vp, I, dp, I]),
        "pb_solver_cta_report": (I, [vp, I, dp, I]),
        "pb_solver_tile_counts": (I, [vp, C.POINTER(LL)]),
        "pb_solver_last_kernel": (C.c_char_p, [vp]),
        "pb_group_create": (vp, [I, I, I, I, C.POINTER(I), I]),
        "pb_group_destroy": (None, [vp]),
        "pb_group_solve": (I, [vp, C.POINTER(SolveOptions), C.POINTER(Stats)]),
        "pb_group_copy_solution": (I, [vp, vp]),
        "pb_step_set_options": (I, [vp, I, D, LL, vp]),
        "pb_step_assemble": (I, [vp, I, vp]),
        "pb_step_begin": (I, [vp, I, vp]),
        "pb_step_dot_zr": (I, [vp, I, vp]),
        "pb_step_stencil": (I, [vp, I, I, I, vp]),
        "pb_step_update": (I, [vp, I, I, vp]),
        "pb_step_update_p": (I, [vp, I, vp]),
        "pb_step_reduce_local": (I, [vp, I, I, vp]),
        "pb_step_scalars": (I, [vp, I, I, vp]),
        "pb_step_pack_edges": (I, [vp, I, vp, I, vp]),
        "pb_step_unpack_edges": (I, [vp, I, vp, I, vp]),
        "pb_step_error_report": (I, [vp, I, vp]),
        "pb_apply_A_dot": (I, [I, I, I, I, I, vp, vp, vp, vp, vp, vp]),
        "pb_apply_A_dot_noalloc": (I, [I, I, I, I, I, vp, vp, vp, vp, vp, vp]),
        "pb_solver_xgpu_bench": (I, [vp, I, I, I, I, vp, dp]),
        "pb_debug_div_check": (I, [D, C.c_ulonglong, C.c_ulonglong, vp, vp]),
        "pb_debug_tma_probe": (I, [vp, I, I, I, I, I, C.c_uint, vp, vp, vp]),
        "pb_step_read_control": (I, [vp, I, C.POINTER(I), C.POINTER(I), C.POINTER(D), vp]),
    })
    for name, (res, args) in sig.items():
        fn = getattr(lib, name, None)
        if fn is None:  # an older library selected through POISSON_B200_LIB (A/B runs): the call itself will fail
            continue
        fn.restype = res
        fn.argtypes = args


def lib() -> C.CDLL:
    """Load (building first if necessary) the native library."""
    global _LIB
    if _LIB is None:
        path = Path(os.environ.get("POISSON_B200_LIB", _build.LIB_PATH))
        if not path.exists() or os.environ.get("POISSON_B200_REBUILD"):
            _build.build_native()
        _LIB = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
        _declare(_LIB)
    return _LIB


def last_error() -> str:
    return lib().pb_last_error().decode()


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {last_error()}")
