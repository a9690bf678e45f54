"""pb_apply_A_dot_noalloc: the operator entry point for callers that are not allowed to allocate.

Collected last on purpose (file name): a failure here must not hide the rest of the suite from a run that
stops at the first failure.  The entry point is driven through the C ABI with ctypes, exactly as a foreign
caller would, and compared with pb_apply_A_dot (which takes its scratch from the stream's memory pool) on the
same inputs.  Two calls are made back to back because all calls share one built-in scratch array that keeps
the previous call's contents: the second call must be as good as the first.  Works both in a fresh process
(it is then the first thing to touch the library on the device, so it also exercises the one-time address
look-up) and after the other tests have used the GPU.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(M, N, seed):
    from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver

    s = GpuSolver(M, N).finalize()
    s.step("assemble")
    torch.cuda.synchronize()
    a, b = s.array("a").clone(), s.array("b").clone()
    blk, co = s.block(), s.col_off
    s.close()
    g = torch.Generator(device="cuda").manual_seed(seed)
    p = torch.zeros_like(a)
    p[1:blk.nx + 1, co + 1:co + 1 + blk.ny] = torch.rand((blk.nx, blk.ny), dtype=torch.float64, device="cuda", generator=g)
    return blk, a, b, p


def _call(fn, M, N, blk, p, a, b):
    Ap = torch.zeros_like(p)
    dot = torch.zeros(1, dtype=torch.float64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = fn(M, N, blk.nx, blk.ny, blk.pitch, C.c_void_p(p.data_ptr()), C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()),
            C.c_void_p(Ap.data_ptr()), C.c_void_p(dot.data_ptr()), stream)
    torch.cuda.synchronize()
    return rc, Ap, dot


def _torch_stencil(M, N, blk, p, a, b):
    """Ap on the owned nodes and h1 h2 (Ap, p), evaluated with plain PyTorch fp64 arithmetic: an oracle that shares
    no code with either entry point."""
    from poisson_ellipse_openmp_mpi_cuda_b200.parallel.decomp import col_offset

    co, h1, h2 = col_offset(), 2.0 / M, 1.2 / N
    I, J = slice(1, blk.nx + 1), slice(co + 1, co + 1 + blk.ny)
    Ip, Im = slice(2, blk.nx + 2), slice(0, blk.nx)
    Jp, Jm = slice(co + 2, co + 2 + blk.ny), slice(co, co + blk.ny)
    ax = -1.0 / h1 * (a[Ip, J] * (p[Ip, J] - p[I, J]) / h1 - a[I, J] * (p[I, J] - p[Im, J]) / h1)
    ay = -1.0 / h2 * (b[I, Jp] * (p[I, Jp] - p[I, J]) / h2 - b[I, J] * (p[I, J] - p[I, Jm]) / h2)
    ref = ax + ay
    return (I, J), ref, float((ref * p[I, J]).sum()) * h1 * h2


def test_noalloc_entry_point_matches_allocating_one(native):
    lib = native.lib()
    torch.cuda.set_device(0)
    # two different problems in a row: the second call finds the first call's partial sums in the shared scratch
    for (M, N, seed) in [(400, 600, 1), (160, 220, 2)]:
        blk, a, b, p = _inputs(M, N, seed)
        rc_ref, Ap_ref, dot_ref = _call(lib.pb_apply_A_dot, M, N, blk, p, a, b)
        assert rc_ref == 0, native.last_error()
        (I, J), Ap_torch, dot_torch = _torch_stencil(M, N, blk, p, a, b)
        for _ in range(2):
            rc, Ap, dot = _call(lib.pb_apply_A_dot_noalloc, M, N, blk, p, a, b)
            assert rc == 0, native.last_error()
            assert torch.allclose(Ap, Ap_ref, rtol=1e-13, atol=1e-12)
            assert dot.item() == pytest.approx(dot_ref.item(), rel=1e-12)
            assert dot.item() != 0.0
            # and against the independent PyTorch evaluation, not only against the sibling entry point (the two
            # differ by roundings of individual terms of size ~ a/h^2, so the bound is relative to the largest
            # value, not to each node's own, possibly cancelled, value)
            assert float((Ap[I, J] - Ap_torch).abs().max()) <= 1e-13 * float(Ap_torch.abs().max())
            assert dot.item() == pytest.approx(dot_torch, rel=1e-12)


def test_noalloc_entry_point_rejects_blocks_beyond_its_scratch(native):
    lib = native.lib()
    # 32768 x 16384 owned nodes -> 1024 * 256 = 262144 tiles > 65536; rejected before anything is dereferenced
    rc = lib.pb_apply_A_dot_noalloc(32769, 16385, 32768, 16384, 16400, None, None, None, None, None, None)
    assert rc != 0
    msg = native.last_error()
    assert "65536" in msg and "tiles" in msg
