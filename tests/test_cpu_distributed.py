"""Distributed CPU solver over gloo (the reference's stage 2 / stage 3 capability), world_size > 1."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ERRORS, ITERS_WEIGHTED


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, M, N, grid, threads, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from poisson_ellipse_openmp_mpi_cuda_b200.parallel.collective import CollectiveSolver

    s = CollectiveSolver(M, N, grid=grid, device="cpu", threads=threads)
    res = s.solve()
    w = s.solution_interior()
    q.put((rank, res, s.blk.i0, s.blk.j0, w))
    dist.barrier()
    dist.destroy_process_group()


def _run(world, M, N, grid=None, threads=1):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, M, N, grid, threads, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    return sorted(out, key=lambda t: t[0])


@pytest.mark.parametrize("world,grid", [(2, None), (4, None), (4, "1x4"), (3, None)])
def test_gloo_matches_oracle(native, world, grid):
    import ctypes as C

    M, N = 40, 40
    out = _run(world, M, N, grid)
    ref = torch.zeros((M + 1, N + 1), dtype=torch.float64)
    st = native.Stats()
    native.lib().pb_cpu_solve(M, N, 0, 1e-6, -1, 1, 0, ref.data_ptr(), C.byref(st))
    for rank, res, i0, j0, w in out:
        assert res["iters"] == ITERS_WEIGHTED[(M, N)] == st.iters
        assert res["status"] == 1
        assert res["l2_in_D"] == pytest.approx(ERRORS[(M, N)][0], rel=1e-7)
        assert res["nodes_in_D"] == 1039
        nx, ny = w.shape
        assert torch.allclose(w, ref[i0:i0 + nx, j0:j0 + ny], rtol=0, atol=1e-12)


def test_gloo_hybrid_threads(native):
    """Stage-3 analogue: 2 ranks x 2 OpenMP threads."""
    out = _run(2, 160, 160, None, threads=2)
    assert all(res["iters"] == 183 for _, res, *_ in out)
