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


@pytest.mark.parametrize("method", ["persistent", "persistent_tma", "resident", "persistent_ldg", "step_fused", "step_classic"])
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


@pytest.mark.parametrize("method", ["persistent_tma", "resident"])
@pytest.mark.parametrize("pg", [(1, 2), (2, 1), (2, 2), (2, 4), (3, 3)])
@pytest.mark.parametrize("grid", [(40, 40), (160, 160), (400, 600)])
def test_virtual_ranks_match_single(grid, pg, method):
    """Px x Py virtual ranks on one GPU (one cooperative launch): same iterations, same solution."""
    one = _solver(*grid)
    r1 = one.solve(method=method)
    w1 = one.solution_host()
    many = _solver(*grid, pg[0], pg[1])
    rm = many.solve(method=method)
    assert rm["kernel"] == ("tma" if method == "persistent_tma" else "resident")
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


@pytest.mark.parametrize("method,kernel", [("persistent", "tma"), ("resident", "resident")])
def test_large_grid_800x1200(method, kernel):
    s = _solver(800, 1200)
    res = s.solve(method=method)
    assert res["iters"] == 989 and res["kernel"] == kernel
    _check_errors(res, (800, 1200))
    s.close()


def test_resident_kernel_is_refused_for_blocks_that_do_not_fit():
    s = _solver(160, 160)
    assert s.solve()["kernel"] == "resident"  # auto: small blocks stay on chip
    s.close()
    s = _solver(2048, 2048)
    assert s.solve(max_iter=5)["kernel"] == "tma"  # auto: too large for shared memory
    with pytest.raises(RuntimeError, match="does not fit"):
        s.solve(method="resident", max_iter=5)
    s.close()


def test_apply_A_dot_operator_matches_fp32_free_reference(native):
    """Stand-alone Ap = A p + (Ap, p): against a plain PyTorch evaluation of the same stencil (fp64) and the CPU
    routine of the oracle."""
    from poisson_ellipse_openmp_mpi_cuda_b200.ops.operators import apply_A_dot

    M, N = 200, 300
    s = _solver(M, N).finalize()
    s.step("assemble")
    torch.cuda.synchronize()
    a, b = s.array("a").clone(), s.array("b").clone()
    blk, co = s.block(), s.col_off
    g = torch.Generator(device="cuda").manual_seed(7)
    p = torch.zeros_like(a)
    p[1:blk.nx + 1, co + 1:co + 1 + blk.ny] = torch.rand((blk.nx, blk.ny), dtype=torch.float64, device="cuda", generator=g)
    Ap, dot = apply_A_dot(M, N, p, a, b)
    h1, h2 = 2.0 / M, 1.2 / N
    I, J = slice(1, blk.nx + 1), slice(co + 1, co + 1 + blk.ny)
    Ip, Im = slice(2, blk.nx + 2), slice(0, blk.nx)
    Jp, Jm = slice(co + 2, co + 2 + blk.ny), slice(co, co + blk.ny)
    ax = -1.0 / h1 * (a[Ip, J] * (p[Ip, J] - p[I, J]) / h1 - a[I, J] * (p[I, J] - p[Im, J]) / h1)
    ay = -1.0 / h2 * (b[I, Jp] * (p[I, Jp] - p[I, J]) / h2 - b[I, J] * (p[I, J] - p[I, Jm]) / h2)
    ref = ax + ay
    assert torch.allclose(Ap[I, J], ref, rtol=1e-13, atol=1e-9)
    assert dot.item() == pytest.approx((ref * p[I, J]).sum().item() * h1 * h2, rel=1e-12)
    # CPU oracle routine on the same data
    nb = native.Block()
    native.lib().pb_local_block(M, N, 1, 1, 0, C.byref(nb))
    ph, ah, bh = p.cpu(), a.cpu(), b.cpu()
    Aph = torch.zeros_like(ph)
    s_cpu = native.lib().pb_cpu_apply_A_local(M, N, C.byref(nb), ph.data_ptr(), ah.data_ptr(), bh.data_ptr(), Aph.data_ptr(), 1)
    assert torch.allclose(Ap.cpu()[I, J], Aph[I, J], rtol=1e-13, atol=1e-9)
    assert dot.item() == pytest.approx(s_cpu * h1 * h2, rel=1e-12)
    s.close()


@pytest.mark.parametrize("pg", [(1, 1), (2, 2), (3, 1)])
@pytest.mark.parametrize("tile_rows", [32, 16])
def test_coefficient_classes_are_bit_identical_to_streamed_coefficients(pg, tile_rows):
    """The 1-byte class path (a, b, D from {1, 1/eps, constants} or a rare array read) must give the same bits as
    streaming the fp64 coefficient boxes."""
    for grid in [(96, 200), (400, 600)]:
        s = _solver(*grid, *pg)
        # (cyclic tile -> CTA map and the tile path of phase AB for both: the balanced map and the warp-private rows
        # depend on the classes, and with them the order of the partial sums)
        ra = s.solve(tile_rows=tile_rows, full_coefficients=True, balance=False, ab_rows=False)
        wa = s.solution_host().clone()
        rb = s.solve(tile_rows=tile_rows, balance=False, ab_rows=False)
        wb = s.solution_host()
        assert ra["iters"] == rb["iters"]
        assert torch.equal(wa, wb)
        s.close()


@pytest.mark.parametrize("pg", [(1, 1), (2, 2), (3, 1)])
@pytest.mark.parametrize("tile_rows", [32, 16])
def test_recomputed_Ap_is_bit_identical_to_stored_Ap(pg, tile_rows):
    """ap_mode=2 (phase C evaluates the stencil again on full constant-coefficient tiles instead of reloading Ap)
    must give the same iteration count, the same stop norm and the same bits as ap_mode=1."""
    for grid in [(96, 200), (400, 600)]:
        s = _solver(*grid, *pg)
        # (cyclic tile -> CTA map for both: the balancer's costs depend on the Ap mode, and with the map the summation order)
        ra = s.solve(tile_rows=tile_rows, ap_mode=1, balance=False)
        wa = s.solution_host().clone()
        rb = s.solve(tile_rows=tile_rows, ap_mode=2, balance=False)
        wb = s.solution_host()
        assert ra["iters"] == rb["iters"] and rb["status"] == 1
        assert ra["diff_norm"] == rb["diff_norm"]
        assert torch.equal(wa, wb)
        s.close()


@pytest.mark.parametrize("method", ["persistent", "persistent_tma"])
def test_per_cta_clocks(method):
    """Every CTA reports where it spent the solve; the four parts add up to (at most) the device time of the loop."""
    s = _solver(800, 1200)
    r = s.solve(method=method)
    t = s.cta_times()
    assert tuple(t.shape) == (r["ctas_per_rank"], 4)
    assert (t >= 0).all() and t[:, :2].min() > 0
    assert float(t.sum(dim=1).max()) <= r["t_iterations"] * 1.05
    s.close()


def test_class_bytes_match_arrays(native):
    s = _solver(200, 300).finalize()
    s.step("assemble")
    torch.cuda.synchronize()
    a, b, D = s.array("a"), s.array("b"), s.array("D")
    inv_eps = 1.0 / max(2.0 / 200, 1.2 / 300) ** 2
    frac_cut = ((a != 1.0) & (a != inv_eps)).double().mean().item()
    assert 0 < frac_cut < 0.02  # only the faces cut by the ellipse carry a blended coefficient
    s.close()
