"""stdout reports in the reference's exact formats.

Stage 4 (S4:1001-1002, S4:900-902, S4:969-978, S4:1027-1033), stages 2/3 (S2:476-477, S2:494-497,
S3:486-487), stage 1 (S1:209-223) and stage 0 (S0:157-158, S0:189-192).  C++ iostream "default float"
formatting is ``%g`` with 6 significant digits; ``std::fixed << setprecision(6)`` is ``%.6f``.
After the reference's own lines we append the error report and the per-iteration cost, which the
reference does not have.
"""
from __future__ import annotations


def _g(x: float) -> str:
    return "%g" % x


def converged_line(iters: int, delta: float, stage0: bool = False) -> str:
    if stage0:
        return f"Converged after {iters} iterations (||w(k+1)-w(k)|| < δ)."
    return f"Converged after {iters} iterations (||w(k+1)-w(k)|| < {_g(delta)})."


def stage4_report(M: int, N: int, nranks: int, res: dict, delta: float, t_init: float, t_solver: float,
                  t_finalize: float = 0.0, extras: bool = True) -> str:
    lines = [f"MPI + CUDA 2D run with {nranks} processes; M={M}, N={N}"]
    if res.get("status", 1) == 1:
        lines.append(converged_line(res["iters"], delta))
    lines += [
        f"   GPU compute time (Ap + D^{{-1}}r, max over ranks) ~ {_g(res.get('t_gpu', 0.0))} s",
        f"   Host<->Device copy time (max over ranks)        ~ {_g(res.get('t_copy', 0.0))} s",
        f"   MPI halo exchange time (max over ranks)         ~ {_g(res.get('t_comm', 0.0))} s",
        f"   Preconditioner CPU part time (max over ranks)   ~ {_g(res.get('t_precond', 0.0))} s",
        f"   Dot products time (max over ranks)              ~ {_g(res.get('t_dot', 0.0))} s",
        f"M={M}, N={N} | Iter={res['iters']} | Total Time={t_init + t_solver + t_finalize:.6f} s",
        f"   Init time (program)      ~ {t_init:.6f} s",
        f"   Solver time (MPI+CUDA)   ~ {t_solver:.6f} s",
        f"   Finalization time        ~ {t_finalize:.6f} s",
    ]
    if extras:
        lines += extra_lines(res)
    return "\n".join(lines)


def extra_lines(res: dict) -> list[str]:
    out = []
    if "l2_in_D" in res:
        out.append(f"   L2 error vs analytic (in D)  ~ {res['l2_in_D']:.9e}")
        out.append(f"   L2 error vs analytic (all)   ~ {res['l2_all']:.9e}")
        out.append(f"   Max error vs analytic      ~ {res['max_abs']:.9e}")
    if res.get("t_iterations") and res.get("iters"):
        out.append(f"   Device time per iteration    ~ {1e6 * res['t_iterations'] / res['iters']:.3f} us")
    return out


def stage23_report(M: int, N: int, nranks: int, res: dict, delta: float, seconds: float, hybrid: bool,
                   extras: bool = True) -> str:
    head = (f"MPI/OpenMP run with {nranks} MPI processes; M={M}, N={N}" if hybrid
            else f"Pure MPI 2D run with {nranks} processes; M={M}, N={N}")
    lines = [head]
    if res.get("status", 1) == 1:
        lines.append(converged_line(res["iters"], delta))
    lines.append(f"M={M}, N={N} | Iter={res['iters']} | Time={seconds:.6f} s")
    if extras:
        lines += extra_lines(res)
    return "\n".join(lines)


def stage0_report(M: int, N: int, res: dict, seconds: float, extras: bool = True) -> str:
    lines = []
    if res.get("status", 1) == 1:
        lines.append(converged_line(res["iters"], 0.0, stage0=True))
    lines.append(f"M={M}, N={N} | Iter={res['iters']} | Time={seconds:.4f} s")
    if extras:
        lines += extra_lines(res)
    return "\n".join(lines)


STAGE1_RULE = "-" * 56


def stage1_banner(M: int, N: int) -> str:
    """The block the OpenMP program prints before its thread sweep (S1:208-210)."""
    return "\n".join(["--- (Variant 9: Ellipse x^2 + 4y^2 < 1, OpenMP Test) ---", f"Grid: M={M}, N={N}", STAGE1_RULE])


def stage1_report(M: int, N: int, threads: int, res: dict, delta: float, seconds: float) -> str:
    """One entry of the thread sweep (S1:192, S1:221-223); like stage 0 it prints the symbol, not delta's value."""
    lines = []
    if res.get("status", 1) == 1:
        lines.append(converged_line(res["iters"], delta, stage0=True))
    lines.append("Threads = %2d | Time = %.3f s" % (threads, seconds))
    return "\n".join(lines)


def stage1_program(M: int, N: int, entries: list) -> str:
    """Whole stdout of the OpenMP program: banner, one block per (threads, res, seconds) entry, closing rule (S1:226)."""
    body = [stage1_report(M, N, t, res, 0.0, sec) for t, res, sec in entries]
    return "\n".join([stage1_banner(M, N), *body, STAGE1_RULE])
