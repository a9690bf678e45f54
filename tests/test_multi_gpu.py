"""Multi-GPU paths (need >= 2 devices): fused peer-memory solver over CUDA IPC, NCCL baseline, in-process group."""
import ctypes as C
import os
import socket

import pytest
import torch

from conftest import ERRORS, ITERS_WEIGHTED

pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, M, N, backend, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import poisson_b200 as pb

    s = pb.PoissonSolver(M, N, backend=backend)
    res = s.solve(watchdog_s=10.0) if backend == "fused" else s.solve()
    res2 = s.solve(watchdog_s=10.0) if backend == "fused" else None
    w = s.solution()
    q.put((rank, {k: v for k, v in res.items() if not isinstance(v, torch.Tensor)}, res2 and res2["iters"], w))
    s.close()
    dist.barrier()
    dist.destroy_process_group()


def _run(world, M, N, backend):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, M, N, backend, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    return sorted(out, key=lambda t: t[0])


def _reference_field(native, M, N):
    ref = torch.zeros((M + 1, N + 1), dtype=torch.float64)
    st = native.Stats()
    native.lib().pb_cpu_solve(M, N, 0, 1e-6, -1, 1, 0, ref.data_ptr(), C.byref(st))
    return ref, st.iters


@pytest.mark.parametrize("backend", ["fused", "nccl"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_process(native, backend, world):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    M, N = 160, 160
    out = _run(world, M, N, backend)
    ref, it = _reference_field(native, M, N)
    for rank, res, it2, w in out:
        assert res["iters"] == ITERS_WEIGHTED[(M, N)] == it
        assert res["status"] == 1
        assert it2 in (None, it)
        assert res["l2_in_D"] == pytest.approx(ERRORS[(M, N)][0], rel=2e-6)
        assert torch.allclose(w, ref, rtol=0, atol=1e-12)


@pytest.mark.parametrize("ngpu", [2, 4, 8])
def test_in_process_group_cli(ngpu):
    """The CLI drives G GPUs from one process (peer access, one persistent kernel per device)."""
    import subprocess

    from poisson_ellipse_openmp_mpi_cuda_b200 import _build

    if _ngpu() < ngpu:
        pytest.skip(f"needs {ngpu} GPUs")
    out = subprocess.run([str(_build.CLI_PATH), "400", "600", "--gpus", str(ngpu)], capture_output=True, text=True,
                         timeout=120)
    assert out.returncode == 0, out.stderr
    assert f"MPI + CUDA 2D run with {ngpu} processes; M=400, N=600" in out.stdout
    assert "Converged after 546 iterations (||w(k+1)-w(k)|| < 1e-06)." in out.stdout
    assert "M=400, N=600 | Iter=546 | Total Time=" in out.stdout


@pytest.mark.parametrize("ngpu", [2, 8])
def test_python_group_backend(native, ngpu):
    """solve(..., gpus=G): several GPUs from one process, no launcher."""
    import poisson_b200 as pb

    if _ngpu() < ngpu:
        pytest.skip(f"needs {ngpu} GPUs")
    res, w = pb.solve(400, 600, gpus=ngpu, return_solution=True)
    assert res["iters"] == 546 and res["status_name"] == "converged"
    assert res["l2_in_D"] == pytest.approx(ERRORS[(400, 600)][0], rel=2e-6)
    ref, _ = _reference_field(native, 400, 600)
    assert torch.allclose(w, ref, rtol=0, atol=1e-11)  # 546 iterations, different summation order
    pb.clear_cache()
