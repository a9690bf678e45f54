"""Library-collective solvers: the same PCG, but with ``torch.distributed`` between separate local steps.

Two flavours share the driver below:

* ``device="cuda"`` -- the NCCL baseline ("the baseline, not the product"): the step-wise sm_100a kernels
  (csrc/cuda/step_kernels.cu) with ``isend/irecv`` halo exchange of ``p`` and ``all_reduce`` of the scalars.
  It has the structure of the reference's stage 4 (S4:846-942) minus the host staging.
* ``device="cpu"`` -- the distributed CPU solver over gloo: the capabilities of the reference's stage 2
  (pure MPI, S2:356-460) and, with ``threads > 1``, stage 3 (MPI+OpenMP, S3:364-470).  Local loops are the
  native OpenMP routines of csrc/cpu/oracle.cpp.
"""
from __future__ import annotations

import ctypes as C
import time

import torch
import torch.distributed as dist

from ..ops import _native
from .decomp import col_offset, local_block, native_block, parse_grid
from .fabric import combine_reports, world_info

_BREAKDOWN = 1e-15


def _exchange(sends: dict, recvs: dict):
    """sends / recvs: {peer_rank: tensor}.  Posts everything, then waits (S2:283-327 style)."""
    ops = []
    for peer, t in recvs.items():
        ops.append(dist.P2POp(dist.irecv, t, peer))
    for peer, t in sends.items():
        ops.append(dist.P2POp(dist.isend, t, peer))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()


class CollectiveSolver:
    def __init__(self, M: int, N: int, grid=None, device: str = "cuda", threads: int = 1, native_cpu: bool = True):
        self.rank, self.world = world_info()
        self.Px, self.Py = parse_grid(grid, self.world)
        self.M, self.N = M, N
        self.kind = device
        self.threads = threads
        self.blk = local_block(M, N, self.Px, self.Py, self.rank)
        self.co = col_offset()
        self.h1, self.h2 = 2.0 / M, 1.2 / N
        self.h1h2 = self.h1 * self.h2
        if device == "cuda":
            from ..ops.gpu import GpuSolver

            # a private single-rank arena; peers are reached through the collective library only
            self.gpu = GpuSolver.__new__(GpuSolver)
            GpuSolver.__init__(self.gpu, M, N, self.Px, self.Py, first_rank=self.rank, nlocal=1)
            self.gpu.finalize()  # peers stay unmapped: no direct stores, halos go through the library
            dev = f"cuda:{self.gpu.device}"
            n = max(self.blk.nx, self.blk.ny)
            self.stride = n
            self.sendbuf = torch.zeros(4 * n, dtype=torch.float64, device=dev)
            self.recvbuf = torch.zeros(4 * n, dtype=torch.float64, device=dev)
            self.red = self.gpu.control_red(0)
        else:
            shape = (self.blk.nx + 2, self.blk.pitch)
            self.f = {k: torch.zeros(shape, dtype=torch.float64) for k in ("w", "r", "p", "Ap", "a", "b", "D")}
            self.nblk = native_block(self.blk)
            # The host loop runs in C++ over the shared-memory transport (csrc/cpu/dist_solver.cpp) when every rank
            # sits on this node; gloo and the Python loop below remain for runs that span nodes.
            self.native_cpu = native_cpu and self._single_node()
            self._jobs = 0

    # ------------------------------------------------------------------------------------------
    def solve(self, norm: str = "weighted", delta: float = 1e-6, max_iter: int = -1, check_every: int = 16,
              div: str = "markstein") -> dict:
        if max_iter < 0:
            max_iter = (self.M - 1) * (self.N - 1)
        if self.kind == "cuda":
            res = self._solve_cuda(norm, delta, max_iter, check_every, div)
            return combine_reports(res, self.M, self.N, device=f"cuda:{self.gpu.device}")
        if self.native_cpu:
            return self._solve_cpu_native(norm, delta, max_iter)
        res = self._solve_cpu(norm, delta, max_iter)
        return combine_reports(res, self.M, self.N, device="cpu")

    # ---- NCCL baseline ----------------------------------------------------------------------
    def _halo_cuda(self):
        g, b, n = self.gpu, self.blk, self.stride
        g.step("pack_edges", self.sendbuf.data_ptr(), n)
        lens = (b.ny, b.ny, b.nx, b.nx)
        sends, recvs = {}, {}
        for d in range(4):
            if b.nbr[d] >= 0:
                sends[b.nbr[d]] = self.sendbuf[d * n:d * n + lens[d]]
                recvs[b.nbr[d]] = self.recvbuf[d * n:d * n + lens[d]]
        _exchange(sends, recvs)
        g.step("unpack_edges", self.recvbuf.data_ptr(), n)

    def _allreduce_cuda(self):
        if self.world > 1:
            dist.all_reduce(self.red[:2])

    def _solve_cuda(self, norm, delta, max_iter, check_every, div):
        from ..ops.gpu import DIVS, NORMS

        g = self.gpu
        ev0, ev1, ev2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        g.step("set_options", NORMS[norm], delta, max_iter)
        ev0.record()
        g.step("assemble")
        ev1.record()
        g.step("begin")
        g.step("dot_zr")
        g.step("reduce_local", 1)
        self._allreduce_cuda()
        g.step("scalars", 0)
        launches = 5
        k = 0
        ctl = {"status": 3}
        while k < max_iter:
            for _ in range(min(check_every, max_iter - k)):
                g.step("update_p")
                self._halo_cuda()
                g.step("stencil", 0, DIVS[div])
                g.step("reduce_local", 1)
                self._allreduce_cuda()
                g.step("scalars", 1)
                g.step("update", 0)
                g.step("reduce_local", 2)
                self._allreduce_cuda()
                g.step("scalars", 2)
                launches += 10
                k += 1
            ctl = g.read_control()
            if ctl["status"] != 3:
                break
        g.step("error_report")
        ev2.record()
        torch.cuda.synchronize()
        ctl = g.read_control()
        err = g.control_err(0).tolist()
        return {
            "iters": ctl["iters"], "status": ctl["status"], "diff_norm": ctl["diff_norm"],
            "sum_e2_in_D": err[0], "sum_e2_all": err[1], "max_abs": err[2], "max_w": err[3], "nodes_in_D": err[4],
            "l2_in_D": (self.h1h2 * err[0]) ** 0.5, "l2_all": (self.h1h2 * err[1]) ** 0.5,
            "seconds": 1e-3 * ev0.elapsed_time(ev2), "t_assemble": 1e-3 * ev0.elapsed_time(ev1),
            "t_iterations": 1e-3 * ev1.elapsed_time(ev2), "t_gpu": 1e-3 * ev1.elapsed_time(ev2), "t_copy": 0.0,
            "t_comm": 0.0, "t_precond": 0.0, "t_dot": 0.0, "launches": launches,
            "status_name": {0: "max_iter", 1: "converged", 2: "breakdown"}.get(ctl["status"], "?"),
        }

    # ---- native CPU solver (C++ host loop, shared-memory transport) ---------------------------
    def _single_node(self) -> bool:
        if self.world == 1:
            return True
        import socket

        hosts = [None] * self.world
        dist.all_gather_object(hosts, socket.gethostname())
        return len(set(hosts)) == 1

    def _solve_cpu_native(self, norm, delta, max_iter):
        import os

        from ..ops.gpu import NORMS, STATUS

        self._jobs += 1
        job = [f"{os.getpid()}_{self._jobs}"]  # rank 0's name for the shared-memory segment of this solve
        if self.world > 1:
            dist.broadcast_object_list(job, src=0)
        st = _native.Stats()
        _native.check(_native.lib().pb_cpu_dist_solve(self.M, self.N, self.Px, self.Py, self.rank, job[0].encode(),
                                                      NORMS[norm], delta, max_iter, self.threads,
                                                      self.f["w"].data_ptr(), C.byref(st)), "pb_cpu_dist_solve")
        res = st.as_dict()
        res["status_name"] = STATUS.get(st.status, "?")
        res["launches"] = 0
        res["host_loop"] = "c++ (shared memory)"
        return res

    # ---- gloo CPU solver --------------------------------------------------------------------
    def _halo_cpu(self):
        p, b, co = self.f["p"], self.blk, self.co
        cols = slice(co + 1, co + 1 + b.ny)
        rows = slice(1, b.nx + 1)
        send = {0: p[1, cols], 1: p[b.nx, cols], 2: p[rows, co + 1], 3: p[rows, co + b.ny]}
        sends, recvs, stage = {}, {}, {}
        for d in range(4):
            if b.nbr[d] >= 0:
                sends[b.nbr[d]] = send[d].contiguous()
                stage[d] = torch.empty_like(sends[b.nbr[d]])
                recvs[b.nbr[d]] = stage[d]
        _exchange(sends, recvs)
        if 0 in stage:
            p[0, cols] = stage[0]
        if 1 in stage:
            p[b.nx + 1, cols] = stage[1]
        if 2 in stage:
            p[rows, co] = stage[2]
        if 3 in stage:
            p[rows, co + b.ny + 1] = stage[3]

    def _allreduce_cpu(self, *vals):
        t = torch.tensor(vals, dtype=torch.float64)
        if self.world > 1:
            dist.all_reduce(t)
        return t.tolist()

    def _solve_cpu(self, norm, delta, max_iter):
        lib, f, nb, T = _native.lib(), self.f, self.nblk, self.threads
        ptr = {k: v.data_ptr() for k, v in f.items()}
        if self.world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        t_comm = 0.0
        for k in ("w", "p", "Ap"):
            f[k].zero_()
        lib.pb_cpu_assemble_local(self.M, self.N, C.byref(nb), ptr["a"], ptr["b"], ptr["D"], ptr["r"], T)
        t_asm = time.perf_counter() - t0
        (zr_old,) = self._allreduce_cpu(lib.pb_cpu_dot_zr_local(C.byref(nb), ptr["r"], ptr["D"], T))
        zr_old *= self.h1h2
        beta, status, iters, diff_norm = 0.0, 0, 0, 0.0
        out2 = (C.c_double * 2)()
        for k in range(1, max_iter + 1):
            iters = k
            lib.pb_cpu_update_p_local(C.byref(nb), ptr["p"], ptr["r"], ptr["D"], beta, T)
            tc = time.perf_counter()
            self._halo_cpu()
            t_comm += time.perf_counter() - tc
            local = lib.pb_cpu_apply_A_local(self.M, self.N, C.byref(nb), ptr["p"], ptr["a"], ptr["b"], ptr["Ap"], T)
            tc = time.perf_counter()
            (denom,) = self._allreduce_cpu(local)
            t_comm += time.perf_counter() - tc
            denom *= self.h1h2
            if abs(denom) < _BREAKDOWN:
                status = 2
                break
            alpha = zr_old / denom
            lib.pb_cpu_update_wr_local(C.byref(nb), ptr["w"], ptr["r"], ptr["p"], ptr["Ap"], ptr["D"], alpha, out2, T)
            tc = time.perf_counter()
            diff2, zr_new = self._allreduce_cpu(out2[0], out2[1])
            t_comm += time.perf_counter() - tc
            zr_new *= self.h1h2
            diff_norm = (diff2 * self.h1h2) ** 0.5 if norm == "weighted" else diff2 ** 0.5
            if diff_norm < delta:
                status = 1
                break
            beta = zr_new / zr_old
            zr_old = zr_new
        seconds = time.perf_counter() - t0
        err = (C.c_double * 5)()
        lib.pb_cpu_error_local(self.M, self.N, C.byref(nb), ptr["w"], err)
        return {
            "iters": iters, "status": status, "diff_norm": diff_norm,
            "sum_e2_in_D": err[0], "sum_e2_all": err[1], "max_abs": err[2], "max_w": err[3], "nodes_in_D": err[4],
            "l2_in_D": (self.h1h2 * err[0]) ** 0.5, "l2_all": (self.h1h2 * err[1]) ** 0.5,
            "seconds": seconds, "t_assemble": t_asm, "t_iterations": seconds - t_asm, "t_gpu": 0.0, "t_copy": 0.0,
            "t_comm": t_comm, "t_precond": 0.0, "t_dot": 0.0, "launches": 0,
            "status_name": {0: "max_iter", 1: "converged", 2: "breakdown"}[status],
        }

    def solution_interior(self) -> torch.Tensor:
        """Owned nodes of ``w`` as a dense ``(nx, ny)`` tensor (CPU for gloo, device for NCCL)."""
        b, co = self.blk, self.co
        if self.kind == "cuda":
            return self.gpu.gather_block(0)
        return self.f["w"][1:b.nx + 1, co + 1:co + 1 + b.ny].clone()

    def close(self):
        if self.kind == "cuda":
            self.gpu.close()
