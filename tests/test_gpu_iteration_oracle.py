"""The persistent TMA kernel against oracles it shares no code with.

* one and two iterations (`max_iter = 1, 2`): w, r, p and Ap as the kernel leaves them in its arrays against a
  plain PyTorch fp64 evaluation of the same recurrence built from the assembled a, b, D, F -- on a grid with
  several tiles per CTA (400 x 600: 475 tiles of 32 x 64 on 148 CTAs, so the stage ring wraps, stages are reused
  and the cross-phase prefetch is exercised) and on 2 x 2 virtual ranks (halo records, ring recomputation);
* the converged solution element by element against the CPU oracle (csrc/cpu/oracle.cpp) for every combination
  of tile height and Ap mode;
* 50 iterations at the headline size (4096 x 4096) against the step kernels.
The reference's own check is "equal iteration counts" (R4 p.9-12); these tests can fail where that one cannot.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _solver(*a, **k):
    from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver

    return GpuSolver(*a, **k)


def _assembled(M, N):
    """a, b, D, F of the whole domain as (M+1) x (N+1) fp64 CUDA tensors indexed by node (i, j)."""
    s = _solver(M, N).finalize()
    s.step("assemble")
    torch.cuda.synchronize()
    co = s.col_off
    out = {k: s.array(k)[:, co:co + N + 1].clone() for k in ("a", "b", "D", "r")}
    s.close()
    return out["a"], out["b"], out["D"], out["r"]


def _apply_A(p, a, b, h1, h2):
    """(A p) on the interior nodes, association of S4:526-535; p is (M+1) x (N+1) with a zero boundary."""
    c = p[1:-1, 1:-1]
    ax = -1.0 / h1 * (a[2:, 1:-1] * (p[2:, 1:-1] - c) / h1 - a[1:-1, 1:-1] * (c - p[:-2, 1:-1]) / h1)
    ay = -1.0 / h2 * (b[1:-1, 2:] * (p[1:-1, 2:] - c) / h2 - b[1:-1, 1:-1] * (c - p[1:-1, :-2]) / h2)
    out = torch.zeros_like(p)
    out[1:-1, 1:-1] = ax + ay
    return out


def _torch_pcg(M, N, iters):
    """`iters` iterations of the recurrence of S4:802-942 in plain PyTorch fp64; returns w, r, p, Ap of the last one."""
    a, b, D, F = _assembled(M, N)
    h1, h2 = 2.0 / M, 1.2 / N
    inner = torch.zeros_like(F, dtype=torch.bool)
    inner[1:-1, 1:-1] = True
    Dsafe = torch.where(inner, D, torch.ones_like(D))
    w = torch.zeros_like(F)
    r = torch.where(inner, F, torch.zeros_like(F))
    z = r / Dsafe
    p = z.clone()
    zr = (z * r).sum() * h1 * h2
    Ap = p_used = None
    for _ in range(iters):
        Ap = _apply_A(p, a, b, h1, h2)
        alpha = zr / ((Ap * p).sum() * h1 * h2)
        w = w + alpha * p
        r = r - alpha * Ap
        z = r / Dsafe
        zr_new = (z * r).sum() * h1 * h2
        p_used = p
        p = z + (zr_new / zr) * p
        zr = zr_new
    return w, r, p_used, Ap


def _close(got, ref, what):
    scale = float(ref.abs().max())
    err = float((got - ref).abs().max())
    assert err <= 1e-13 * scale + 1e-300, f"{what}: max |diff| {err:.3e} against scale {scale:.3e}"


@pytest.mark.parametrize("iters", [1, 2])
@pytest.mark.parametrize("pg", [(1, 1), (2, 2)])
@pytest.mark.parametrize("tile_rows,ap_mode,ab_rows", [(32, 1, True), (32, 2, True), (16, 1, True), (64, 2, True), (64, 1, True),
                                                       (32, 1, False), (64, 2, False)])
def test_first_iterations_against_pytorch(iters, pg, tile_rows, ap_mode, ab_rows):
    """ab_rows: phase AB of full constant-coefficient items by warp-private rows (the default) or on the tile path."""
    M, N = 400, 600
    w_ref, r_ref, p_ref, Ap_ref = _torch_pcg(M, N, iters)
    s = _solver(M, N, *pg)
    res = s.solve(method="persistent_tma", max_iter=iters, tile_rows=tile_rows, ap_mode=ap_mode, ab_rows=ab_rows, error_report=False)
    assert res["iters"] == iters and res["status_name"] == "max_iter"
    pname = "p1" if iters % 2 else "p0"  # the buffer phase AB of the last iteration wrote
    for l in range(pg[0] * pg[1]):
        blk = s.block(l)
        I, J = slice(blk.i0, blk.i0 + blk.nx), slice(blk.j0, blk.j0 + blk.ny)
        _close(s.interior("w", l), w_ref[I, J], f"w block {l}")
        _close(s.interior("r", l), r_ref[I, J], f"r block {l}")
        _close(s.interior(pname, l), p_ref[I, J], f"p block {l}")
        if ap_mode == 1:  # with ap_mode 2 full constant-coefficient tiles never store Ap
            _close(s.interior("Ap", l), Ap_ref[I, J], f"Ap block {l}")
    s.close()


@pytest.mark.parametrize("iters", [1, 2, 3])
@pytest.mark.parametrize("pg", [(1, 1), (2, 2), (1, 3)])
def test_resident_kernel_first_iterations_against_pytorch(iters, pg):
    """The shared-memory-resident kernel (sub-block rims through the global r array, halo records between ranks)
    against the same PyTorch recurrence; it always leaves its direction in p1."""
    M, N = 400, 600
    w_ref, r_ref, p_ref, _ = _torch_pcg(M, N, iters)
    s = _solver(M, N, *pg)
    res = s.solve(method="resident", max_iter=iters, error_report=False)
    assert res["iters"] == iters and res["status_name"] == "max_iter" and res["kernel"] == "resident"
    for l in range(pg[0] * pg[1]):
        blk = s.block(l)
        I, J = slice(blk.i0, blk.i0 + blk.nx), slice(blk.j0, blk.j0 + blk.ny)
        _close(s.interior("w", l), w_ref[I, J], f"w block {l}")
        _close(s.interior("r", l), r_ref[I, J], f"r block {l}")
        _close(s.interior("p1", l), p_ref[I, J], f"p block {l}")
    s.close()


def test_resident_kernel_converged_solution_against_cpu_oracle(native):
    M, N = 400, 600
    ref = torch.zeros((M + 1, N + 1), dtype=torch.float64)
    st = native.Stats()
    native.lib().pb_cpu_solve(M, N, 0, 1e-6, -1, 4, 0, ref.data_ptr(), C.byref(st))
    for pg in [(1, 1), (2, 2), (4, 1)]:
        s = _solver(M, N, *pg)
        res = s.solve(method="resident")
        assert res["iters"] == 546 == st.iters and res["kernel"] == "resident"
        assert float((s.solution_host() - ref).abs().max()) < 1e-10
        s.close()


@pytest.mark.parametrize("tile_rows", [32, 16, 64])
@pytest.mark.parametrize("ap_mode", [1, 2])
def test_converged_solution_elementwise_against_cpu_oracle(native, tile_rows, ap_mode):
    M, N = 400, 600
    ref = torch.zeros((M + 1, N + 1), dtype=torch.float64)
    st = native.Stats()
    native.lib().pb_cpu_solve(M, N, 0, 1e-6, -1, 4, 0, ref.data_ptr(), C.byref(st))
    assert st.iters == 546
    for pg in [(1, 1), (2, 2)]:
        s = _solver(M, N, *pg)
        res = s.solve(tile_rows=tile_rows, ap_mode=ap_mode)
        assert res["iters"] == 546
        w = s.solution_host()
        assert float((w - ref).abs().max()) < 1e-10
        s.close()


@pytest.mark.parametrize("pg", [(1, 1), (2, 2), (1, 3)])
@pytest.mark.parametrize("tile_rows", [32, 64])
def test_phase_ab_by_rows_gives_the_bits_of_the_tile_path(pg, tile_rows):
    """Phase AB by warp-private rows evaluates the same expressions on the same operands as the tile path: after one
    iteration p_new (ring cells of the block included) and Ap are equal bit for bit; w and r carry alpha, whose
    reduction sums the per-thread partials in another order, so they -- and the converged solutions -- agree to
    rounding."""
    M, N = 400, 600
    s = _solver(M, N, *pg)
    kw = dict(method="persistent_tma", tile_rows=tile_rows, ap_mode=1, error_report=False)
    s.solve(max_iter=1, ab_rows=False, **kw)
    ref = [{k: s.array(k, l).clone() for k in ("p1", "Ap", "w", "r")} for l in range(pg[0] * pg[1])]
    s.solve(max_iter=1, ab_rows=True, **kw)
    for l in range(pg[0] * pg[1]):
        assert torch.equal(s.array("p1", l), ref[l]["p1"]), f"p_new block {l}"
        assert torch.equal(s.array("Ap", l), ref[l]["Ap"]), f"Ap block {l}"
        for k in ("w", "r"):
            assert float((s.array(k, l) - ref[l][k]).abs().max()) <= 1e-13 * float(ref[l][k].abs().max()), k
    ra = s.solve(ab_rows=False, **kw)
    wa = s.solution_host().clone()
    rb = s.solve(ab_rows=True, **kw)
    assert ra["iters"] == rb["iters"] == 546
    assert float((s.solution_host() - wa).abs().max()) < 1e-11
    s.close()


@pytest.mark.parametrize("ab_rows", [True, False])
@pytest.mark.parametrize("grid,pg", [((257, 193), (1, 1)), ((513, 385), (2, 2)), ((513, 193), (2, 1)), ((257, 385), (1, 2))])
def test_blocks_whose_edges_are_full_tiles(grid, pg, ab_rows):
    """Blocks of 256 x 192 nodes: every edge of the block is an edge of full 64 x 64 items, so the rows path stores the
    block's ring cells of p_new from its own registers (top / bottom by the first / last warp, left / right by lanes
    0 / 31) and the neighbours' halo rows meet full tiles.  Three iterations against PyTorch: the ring of p is read in
    iterations 2 and 3."""
    M, N = grid
    for tile_rows in (64, 32):
        for iters in (1, 3):
            w_ref, r_ref, p_ref, _ = _torch_pcg(M, N, iters)
            s = _solver(M, N, *pg)
            res = s.solve(method="persistent_tma", max_iter=iters, tile_rows=tile_rows, ap_mode=2, ab_rows=ab_rows, error_report=False)
            assert res["iters"] == iters
            pname = "p1" if iters % 2 else "p0"
            for l in range(pg[0] * pg[1]):
                blk = s.block(l)
                assert blk.nx == 256 and blk.ny == 192
                I, J = slice(blk.i0, blk.i0 + blk.nx), slice(blk.j0, blk.j0 + blk.ny)
                _close(s.interior("w", l), w_ref[I, J], f"w block {l}")
                _close(s.interior("r", l), r_ref[I, J], f"r block {l}")
                _close(s.interior(pname, l), p_ref[I, J], f"p block {l}")
                # the ring of p next to a neighbour holds the neighbour's p (recomputed here from its r): compare it too
                full = s.array(pname, l)
                co = s.col_off
                if blk.i0 > 1:
                    _close(full[0, co + 1:co + 1 + blk.ny], p_ref[blk.i0 - 1, J], f"top ring of p, block {l}")
                if blk.i0 + blk.nx < M:
                    _close(full[blk.nx + 1, co + 1:co + 1 + blk.ny], p_ref[blk.i0 + blk.nx, J], f"bottom ring of p, block {l}")
                if blk.j0 > 1:
                    _close(full[1:blk.nx + 1, co], p_ref[I, blk.j0 - 1], f"left ring of p, block {l}")
                if blk.j0 + blk.ny < N:
                    _close(full[1:blk.nx + 1, co + 1 + blk.ny], p_ref[I, blk.j0 + blk.ny], f"right ring of p, block {l}")
            s.close()


def test_tall_items_are_used_and_counted():
    """tile_rows=64 merges pairs of full constant-coefficient tiles; the kernel's own tile counts must still add up to
    the number of 32 x 64 tiles of the block, and the result must agree with 32-row tiles to rounding."""
    M, N = 1600, 1200
    s = _solver(M, N)
    ra = s.solve(tile_rows=32, ap_mode=2)
    ta = s.tile_counts()
    wa = s.solution_host().clone()
    rb = s.solve(tile_rows=64, ap_mode=2)
    tb = s.tile_counts()
    wb = s.solution_host()
    ntiles = ((M - 1 + 31) // 32) * ((N - 1 + 63) // 64)
    assert sum(ta.values()) == ntiles == sum(tb.values())
    assert ra["iters"] == rb["iters"]
    assert float((wa - wb).abs().max()) < 1e-11
    s.close()


def test_headline_grid_50_iterations_against_step_kernels():
    """4096 x 4096 (the benchmark's grid): 50 iterations of the persistent kernel and of the step kernels (one
    launch per phase, plain loads, grid-wide reductions by separate kernels) leave the same w and stop norm."""
    M = N = 4096
    s = _solver(M, N)
    ra = s.solve(max_iter=50, error_report=False)
    wa = s.interior("w").clone()
    rb = s.solve(method="step_fused", max_iter=50, error_report=False)
    wb = s.interior("w")
    assert ra["iters"] == rb["iters"] == 50
    assert ra["diff_norm"] == pytest.approx(rb["diff_norm"], rel=1e-9)
    assert float((wa - wb).abs().max()) <= 1e-12 * float(wb.abs().max())
    s.close()


def test_soak_200_solves_on_virtual_ranks():
    """200 solves in a row on 2 x 2 virtual ranks (halo records, proxy fences, record-star reductions with their
    sequence numbers running on across launches): every one must give the same iteration count and the same bits.
    (compute-sanitizer is closed on the GPU pool -- profiles/r2/sanitizer/ -- so protocol slips are hunted with
    repetition and oracles instead.)"""
    s = _solver(160, 160, 2, 2)
    first = None
    for i in range(200):
        r = s.solve(method="persistent_tma" if i % 2 else "resident", error_report=False, watchdog_s=5.0)
        assert r["iters"] == 183 and r["status_name"] == "converged", (i, r)
        w = s.solution_host()
        if i < 2:
            first = (first or []) + [w.clone()]  # one per kernel: each must reproduce its own bits
        else:
            assert torch.equal(w, first[i % 2]), i
    s.close()
