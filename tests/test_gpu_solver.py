"""GPU solver against the CPU oracle and the reference's iteration / error tables."""
import ctypes as C

import pytest
import torch

from conftest import ERRORS, ITERS_UNWEIGHTED, ITERS_WEIGHTED

pytestmark = pytest.mark.gpu


def _solver(*a, **k):
    from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver

    return GpuSolver(*a, **k)


def _check_errors(res, key, rel=2e-6):
    l2d, l2a, mx, mw = ERRORS[key]
    assert res["l2_in_D"] == pytest.approx(l2d, rel=rel)
    assert res["l2_all"] == pytest.approx(l2a, rel=rel)
    assert res["max_abs"] == pytest.approx(mx, rel=rel)
    assert res["max_w"] == pytest.approx(mw, rel=rel)


def test_assembly_bit_exact(native):
    """K0 arrays equal the CPU oracle bit for bit (same individually rounded expressions)."""
    M, N = 96, 72
    s = _solver(M, N, 2, 2)
    s.finalize()
    for l in range(4):
        s.step("assemble", local_idx=l)
    torch.cuda.synchronize()
    lib = native.lib()
    for l in range(4):
        blk = native.Block()
        lib.pb_local_block(M, N, 2, 2, l, C.byref(blk))
        shape = (blk.nx + 2, blk.pitch)
        ref = {k: torch.zeros(shape, dtype=torch.float64) for k in "abDr"}
        lib.pb_cpu_assemble_local(M, N, C.byref(blk), ref["a"].data_ptr(), ref["b"].data_ptr(), ref["D"].data_ptr(),
                                  ref["r"].data_ptr(), 1)
        co = s.col_off
        for k in "abDr":
            got = s.array(k, l).cpu()[:, co:co + blk.ny + 2]
            assert torch.equal(got, ref[k][:, co:co + blk.ny + 2]), k
    s.close()


@pytest.mark.parametrize("method", ["persistent", "persistent_ldg", "step_fused", "step_classic"])
@pytest.mark.parametrize("grid", [(10, 10), (40, 40), (160, 160), (400, 600)])
def test_iteration_parity_single_rank(grid, method):
    s = _solver(*grid)
    res = s.solve(method=method)
    assert res["status_name"] == "converged"
    assert res["iters"] == ITERS_WEIGHTED[grid]
    _check_errors(res, grid)
    s.close()


@pytest.mark.parametrize("grid", [(10, 10), (20, 20), (40, 40)])
def test_unweighted_norm(grid):
    s = _solver(*grid)
    res = s.solve(norm="unweighted")
    assert res["iters"] == ITERS_UNWEIGHTED[grid]
    s.close()


@pytest.mark.parametrize("pg", [(1, 2), (2, 1), (2, 2), (2, 4), (3, 3)])
@pytest.mark.parametrize("grid", [(40, 40), (160, 160), (400, 600)])
def test_virtual_ranks_match_single(grid, pg):
    """Px x Py virtual ranks on one GPU (one cooperative launch): same iterations, same solution."""
    one = _solver(*grid)
    r1 = one.solve()
    w1 = one.solution_host()
    many = _solver(*grid, pg[0], pg[1])
    rm = many.solve()
    wm = many.solution_host()
    assert rm["iters"] == r1["iters"] == ITERS_WEIGHTED[grid]
    assert torch.allclose(w1, wm, rtol=0, atol=1e-12)
    _check_errors(rm, grid)
    one.close()
    many.close()


@pytest.mark.parametrize("tile_rows", [16, 32])
@pytest.mark.parametrize("pg", [(1, 1), (2, 2)])
def test_tma_tile_heights(tile_rows, pg):
    for grid in [(40, 40), (400, 600)]:
        s = _solver(*grid, *pg)
        res = s.solve(tile_rows=tile_rows)
        assert res["iters"] == ITERS_WEIGHTED[grid]
        _check_errors(res, grid)
        s.close()


@pytest.mark.parametrize("pg", [(1, 2), (2, 2), (2, 4)])
def test_virtual_ranks_ldg_kernel(pg):
    s = _solver(160, 160, *pg)
    res = s.solve(method="persistent_ldg")
    assert res["iters"] == 183
    s.close()


@pytest.mark.parametrize("div", ["true", "recip", "markstein"])
def test_division_modes_keep_iteration_count(div):
    for grid in [(40, 40), (400, 600)]:
        s = _solver(*grid)
        res = s.solve(div=div)
        assert res["iters"] == ITERS_WEIGHTED[grid]
        s.close()


@pytest.mark.parametrize("method", ["persistent", "step_fused"])
def test_markstein_division_is_bit_identical_to_true_division(method):
    """q' = fma(fma(-q, h, x), y, q), y = RN(1/h), must reproduce x / h exactly: the whole solution agrees bitwise."""
    for grid in [(80, 80), (400, 600), (333, 517)]:
        s = _solver(*grid)
        ra = s.solve(method=method, div="true")
        wa = s.solution_host().clone()
        rb = s.solve(method=method, div="markstein")
        wb = s.solution_host()
        assert ra["iters"] == rb["iters"]
        assert torch.equal(wa, wb)
        s.close()


def test_solution_matches_cpu_oracle(native):
    M, N = 80, 80
    s = _solver(M, N)
    s.solve()
    w = s.solution_host()
    ref = torch.zeros((M + 1, N + 1), dtype=torch.float64)
    st = native.Stats()
    native.lib().pb_cpu_solve(M, N, 0, 1e-6, -1, 1, 0, ref.data_ptr(), C.byref(st))
    assert st.iters == 96
    assert torch.allclose(w, ref, rtol=0, atol=1e-12)
    s.close()


def test_repeated_solves_and_max_iter():
    s = _solver(40, 40, 1, 2)
    a = s.solve()
    b = s.solve()
    assert a["iters"] == b["iters"] == 50
    c = s.solve(max_iter=7)
    assert c["iters"] == 7 and c["status_name"] == "max_iter"
    s.close()


def test_large_grid_800x1200():
    s = _solver(800, 1200)
    res = s.solve()
    assert res["iters"] == 989
    _check_errors(res, (800, 1200))
    s.close()
