#!/usr/bin/env python
"""Communication micro-benchmarks (BASELINE.json config 5): halo exchange and scalar all-reduce, edge 64..16384.

    python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1 bench/halo_allreduce_sweep.py

"ours"  = the in-kernel primitives of the persistent solver, timed on the device (%globaltimer in CTA 0, max
          over ranks): cross-rank reduction of two doubles (one-shot LL records over peer memory), and a
          complete halo exchange = peer stores of `edge` doubles per neighbour + the reduction that publishes them.
"nccl"  = torch.distributed over NCCL on the same GPUs: all_reduce of two doubles; isend/irecv of `edge`
          doubles with every neighbour of the Px x Py grid (CUDA events, max over ranks).
The reference's own path (cudaMemcpy D2H -> MPI_Sendrecv -> H2D, MPI_Allreduce; S4:331-500, S4:871) is measured
by its timers in bench.py --impl reference (T_comm + T_copy per iteration).
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from poisson_ellipse_openmp_mpi_cuda_b200.parallel.fabric import FusedMultiGpuSolver, init_distributed  # noqa: E402


def max_over_ranks(x: float, dev) -> float:
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def main():
    rank, world, local = init_distributed("nccl")
    dev = torch.device("cuda", local)
    iters = int(os.environ.get("SWEEP_ITERS", "2000"))
    G = 16384 + 1
    fs = FusedMultiGpuSolver(G * (2 if world > 1 else 1), G * (4 if world > 2 else 2 if world > 1 else 1))
    blk = fs.block()
    rows = []
    fs.solver.xgpu_bench(50, 0, 0)  # warm-up
    ns, chk = fs.solver.xgpu_bench(iters, 0, 0)
    rows.append({"what": "allreduce_2xf64", "impl": "ours_counter+aggregator (plain-load kernel)",
                 "us": max_over_ranks(ns, dev) / 1e3, "checksum_ok": chk == iters * world * 148})
    fs.solver.xgpu_bench(50, 3, 0)
    ns, chk = fs.solver.xgpu_bench(iters, 3, 0)
    rows.append({"what": "allreduce_2xf64", "impl": "ours_record_star (TMA kernel)", "us": max_over_ranks(ns, dev) / 1e3,
                 "checksum_ok": chk == iters * world * 148})
    nbrs = [n for n in blk.nbr if n >= 0]
    for edge in (64, 256, 1024, 4096, 16384):
        fs.solver.xgpu_bench(20, 1, edge)
        ns, _ = fs.solver.xgpu_bench(iters, 1, edge)
        rows.append({"what": "halo_exchange", "edge_doubles": edge, "impl": "ours_peer_store+sys_fenced_reduction",
                     "us": max_over_ranks(ns, dev) / 1e3, "neighbours": len(nbrs)})
        fs.solver.xgpu_bench(20, 2, edge)
        ns, _ = fs.solver.xgpu_bench(iters, 2, edge)
        rows.append({"what": "halo_exchange", "edge_doubles": edge, "impl": "ours_peer_store+counter+reduction",
                     "us": max_over_ranks(ns, dev) / 1e3, "neighbours": len(nbrs)})
        fs.solver.xgpu_bench(20, 4, edge)
        ns, chk = fs.solver.xgpu_bench(iters, 4, edge)
        rows.append({"what": "halo_exchange", "edge_doubles": edge, "impl": "ours_ll_records+record_star (TMA kernel)",
                     "us": max_over_ranks(ns, dev) / 1e3, "neighbours": len(nbrs),
                     "checksum_ok": chk == iters * world * 148})
    if world > 1:
        x = torch.ones(2, dtype=torch.float64, device=dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(20):
            dist.all_reduce(x)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            dist.all_reduce(x)
        e1.record()
        torch.cuda.synchronize()
        rows.append({"what": "allreduce_2xf64", "impl": "nccl", "us": max_over_ranks(1e3 * e0.elapsed_time(e1) / iters, dev)})
        for edge in (64, 256, 1024, 4096, 16384):
            send = {n: torch.ones(edge, dtype=torch.float64, device=dev) for n in nbrs}
            recv = {n: torch.empty(edge, dtype=torch.float64, device=dev) for n in nbrs}

            def once():
                ops = [dist.P2POp(dist.irecv, recv[n], n) for n in nbrs] + [dist.P2POp(dist.isend, send[n], n) for n in nbrs]
                for r in dist.batch_isend_irecv(ops):
                    r.wait()

            for _ in range(10):
                once()
            torch.cuda.synchronize()
            n_it = max(100, iters // 10)
            e0.record()
            for _ in range(n_it):
                once()
            e1.record()
            torch.cuda.synchronize()
            rows.append({"what": "halo_exchange", "edge_doubles": edge, "impl": "nccl_send_recv",
                         "us": max_over_ranks(1e3 * e0.elapsed_time(e1) / n_it, dev), "neighbours": len(nbrs)})
    if rank == 0:
        for r in rows:
            r.update(n_gpus=world, grid=f"{fs.Px}x{fs.Py}")
            print(json.dumps(r), flush=True)
    fs.close()
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
