"""CPU oracle against the reference's iteration / error tables."""
import ctypes as C

import pytest

from conftest import ERRORS, ITERS_UNWEIGHTED, ITERS_WEIGHTED


def _solve(native, M, N, norm=0, threads=1, signed=0):
    st = native.Stats()
    rc = native.lib().pb_cpu_solve(M, N, norm, 1e-6, -1, threads, signed, None, C.byref(st))
    assert rc == 0
    return st


@pytest.mark.parametrize("grid", [g for g in ITERS_WEIGHTED if g[0] <= 400])
def test_weighted_iterations_and_errors(native, grid):
    st = _solve(native, *grid)
    assert st.status == 1
    assert st.iters == ITERS_WEIGHTED[grid]
    l2d, l2a, mx, mw = ERRORS[grid]
    assert st.l2_in_D == pytest.approx(l2d, rel=1e-9)
    assert st.l2_all == pytest.approx(l2a, rel=1e-9)
    assert st.max_abs == pytest.approx(mx, rel=1e-9)
    assert st.max_w == pytest.approx(mw, rel=1e-9)


@pytest.mark.parametrize("grid", list(ITERS_UNWEIGHTED))
def test_unweighted_iterations_stage0(native, grid):
    st = _solve(native, *grid, norm=1, signed=1)
    assert st.iters == ITERS_UNWEIGHTED[grid]


@pytest.mark.parametrize("threads", [2, 4])
def test_openmp_same_iterations(native, threads):
    assert _solve(native, 40, 40, threads=threads).iters == 50
    assert _solve(native, 160, 160, threads=threads).iters == 183


def test_nodes_in_domain(native):
    assert _solve(native, 40, 40).nodes_in_D == 1039
    assert _solve(native, 10, 10).nodes_in_D == 65
