#!/usr/bin/env python
"""Headline benchmark: PCG time-to-converge and us/iteration on the 4096 x 4096 grid (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # ours (N > 1: launched by torchrun)
    python bench.py --impl reference --gpus N ...            # the unmodified reference (baseline/_ref)

One "step" is one complete solve: device-side assembly of a, b, D, F, the PCG iteration to
||w(k+1)-w(k)|| < 1e-6 (3226 iterations at 4096^2) and the error report.  Strong scaling: the grid is fixed,
the N GPUs form the reference's Px x Py block grid.  Rank 0 prints one JSON line.

Timing: every solve is timed on the device with CUDA events (inside the solver, around assembly +
iteration + error report), the max over ranks is taken per solve, `value` is the mean over the K timed
solves.  Between solves a 512 MiB buffer is written (L2 flush); at N = 1 the working set (1.07 GB) exceeds
L2 anyway.  `e2e` is the wall-clock time of the public API call `poisson_b200.solve(M, N)`, which every step
copies the solve options host->device from pinned memory and the result block device->host.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

M_DEFAULT, N_DEFAULT = 4096, 4096
# what the reference published (R4 p.11 Table 1, Tesla P100): whole-solve seconds by
# (M, N) and GPU count.  It published nothing for 4096 x 4096, so `vs_baseline` is null there; the same-box
# comparison is the driver's own: it runs `--impl reference` next to this arm.
BASELINE_MD_PUBLISHED_S = {(800, 1200): {1: 0.83, 2: 0.64}, (1600, 2400): {1: 4.85, 2: 3.19},
                           (2400, 3200): {1: 13.24, 2: 7.67}}
EXPECTED_ITERS = {(4096, 4096): 3226, (800, 1200): 989, (40, 40): 50, (400, 600): 546}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU every 200 ms through NVML."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.nv = None

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def finish(self) -> dict:
        self._stop_evt.set()
        if self.is_alive():
            self.join(2.0)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def measured_hbm_gbs() -> tuple[float, str]:
    """Copy bandwidth of this pool's B200 from the driver-written MEASURED_PEAKS.json (else the profiling recipe's
    fallback, and say so)."""
    try:
        return float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


# In-kernel all-reduce of two doubles, one reduction, microseconds by GPU count, as measured with
# bench/halo_allreduce_sweep.py on this pool (profiles/r2/campaign_*gpu/halo_allreduce_sweep.jsonl; 1 GPU: record star,
# several GPUs: counter + elected aggregator).  The latency floor of an iteration is two of them.
REDUCTION_FLOOR_US = {1: 3.3, 2: 5.5, 4: 6.0, 8: 6.4}
NVLINK_GBS_PER_DIRECTION = 900.0


def roofline(M: int, N: int, world: int, grid, tiles: dict, us_iter: float, l2_bytes: int = 126 << 20) -> dict:
    """The north-star roofline of one iteration on the busiest GPU: max(HBM bytes at the copy bandwidth, halo bytes
    over NVLink) + the latency floor of the two reductions.  Bytes are counted from the kernel's own tile classes
    (64 / 80 / 112 B per node and iteration), not assumed."""
    nodes = (M - 1) * (N - 1)
    nt = max(1, sum(tiles.values()))
    bytes_node = (64.0 * tiles["recompute_64B"] + 80.0 * tiles["stored_80B"] + 112.0 * tiles["mixed_112B"]) / nt
    bytes_iter = bytes_node * nodes / world
    hbm_gbs, src = measured_hbm_gbs()
    px, py = grid
    # halo: every shared edge node travels as one 16-byte record per iteration, both directions at once
    edge_nodes = (min(2, px - 1) * ((N - 1) // py + 1) if px > 1 else 0) + (min(2, py - 1) * ((M - 1) // px + 1) if py > 1 else 0)
    nvlink_s = edge_nodes * 16.0 / (NVLINK_GBS_PER_DIRECTION * 1e9)
    hbm_s, hbm_s_nominal = bytes_iter / (hbm_gbs * 1e9), bytes_iter / 8e12
    red_s = 2e-6 * REDUCTION_FLOOR_US.get(world, REDUCTION_FLOOR_US[8])
    model_s = max(hbm_s, nvlink_s) + red_s
    working_set = 5 * 8 * nodes / world
    return {"bytes_per_node_iter": bytes_node, "tiles": tiles, "bytes_per_iter_per_gpu": bytes_iter,
            "hbm_gbs_used_for_roofline": hbm_gbs, "hbm_gbs_source": src,
            "hbm_us": 1e6 * hbm_s, "hbm_us_at_8TBs": 1e6 * hbm_s_nominal, "nvlink_us": 1e6 * nvlink_s,
            "reduction_floor_us": 1e6 * red_s, "model_us": 1e6 * model_s, "measured_us": us_iter,
            "fraction_of_roofline": model_s / (us_iter * 1e-6) if us_iter > 0 else None,
            "fraction_of_hbm_bandwidth": hbm_s / (us_iter * 1e-6) if us_iter > 0 else None,
            "l2_resident": working_set < 0.8 * l2_bytes,
            "note": "an L2-resident block (w, r, p, p', Ap of one GPU < 80 % of L2) is not bound by HBM: its fraction "
                    "says how close the iteration is to what HBM would allow, not to a roof it could hit"}


def cli_cold_seconds(M: int, N: int, gpus: int) -> dict | None:
    """Process wall time and printed lines of the product's CLI -- the reference's only interface -- from a cold
    start: `poisson_b200 M N --gpus G` (one process drives all G GPUs)."""
    exe = ROOT / "poisson_ellipse_openmp_mpi_cuda_b200" / "bin" / "poisson_b200"
    if not exe.exists():
        return None
    t0 = time.perf_counter()
    p = subprocess.run([str(exe), str(M), str(N), "--gpus", str(gpus)], capture_output=True, text=True)
    wall = time.perf_counter() - t0
    if p.returncode != 0:
        return {"error": p.stderr[-200:], "process_wall_s": wall}
    return {"process_wall_s": wall, "printed_total_time_s": _grab(r"Total Time=([0-9.eE+-]+) s", p.stdout),
            "printed_solver_time_s": _grab(r"Solver time \(MPI\+CUDA\)\s+~ ([0-9.eE+-]+)", p.stdout),
            "printed_init_time_s": _grab(r"Init time \(program\)\s+~ ([0-9.eE+-]+)", p.stdout)}


def _env():
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0)))


# --------------------------------------------------------------------------------------------------
def run_ours(a) -> dict:
    import torch
    import torch.distributed as dist

    import poisson_b200 as pb
    from poisson_ellipse_openmp_mpi_cuda_b200.parallel.fabric import init_distributed

    rank, world, local = init_distributed("nccl")
    if world != a.gpus:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {a.gpus}")
    dev = torch.device("cuda", local)
    M, N = a.M, a.N
    # one solver for both measurements: the object the public pb.solve(M, N, ...) creates and caches
    solver = pb.get_solver(M, N, backend="fused", grid=a.grid)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    kw = dict(div=a.div, tile_rows=a.tile_rows, ap_mode=a.ap_mode)
    for _ in range(a.warmup):
        flush.fill_(1)
        res = solver.solve(**kw)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    # ---- device-timed steps ----
    sync()
    t_wall0 = time.perf_counter()
    dev_s, iters, launches = [], [], 0
    for _ in range(a.steps):
        flush.fill_(1)
        res = solver.solve(**kw)  # 'seconds' = CUDA-event time, max over ranks
        dev_s.append(res["seconds"])
        iters.append(res["iters"])
        launches += res.get("launches", 0)
    sync()
    t_wall = time.perf_counter() - t_wall0
    # ---- end-to-end through the public API (same cached solver; one more untimed call first) ----
    e2e_s = []
    pb.solve(M, N, backend="fused", grid=a.grid, **kw)
    for _ in range(a.steps):
        flush.fill_(1)
        sync()
        t0 = time.perf_counter()
        r2 = pb.solve(M, N, backend="fused", grid=a.grid, **kw)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e_s.append(dt.item())
    clocks = sampler.finish() if sampler else None
    gpu_solver = solver._impl.solver if world > 1 else solver._impl
    info = gpu_solver.info()
    recomputes = gpu_solver.recomputes_ap(a.ap_mode)
    # tile classes counted by the kernel, summed over the ranks -> bytes actually moved per node and iteration
    tc = gpu_solver.tile_counts()
    tct = torch.tensor([tc["recompute_64B"], tc["stored_80B"], tc["mixed_112B"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tct)
    tiles = dict(zip(("recompute_64B", "stored_80B", "mixed_112B"), (int(v) for v in tct.tolist())))
    value = sum(dev_s) / len(dev_s)
    e2e = sum(e2e_s) / len(e2e_s)
    us_iter = 1e6 * res["t_iterations"] / max(1, res["iters"])
    nodes = (M - 1) * (N - 1)
    roof = roofline(M, N, world, solver.grid, tiles, us_iter)
    bytes_node = roof["bytes_per_node_iter"]
    out = {
        "metric": f"pcg_time_to_converge_{M}x{N}", "value": value, "unit": "s", "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * value, "higher_is_better": False, "scaling": "strong",
        "vs_baseline": (value / BASELINE_MD_PUBLISHED_S[(M, N)][world]) if world in BASELINE_MD_PUBLISHED_S.get((M, N), {})
        else None, "dtype": "fp64", "data": "synthetic (analytic problem: coefficients and RHS assembled "
        "on the device from the ellipse geometry; no dataset exists for this solver)", "impl": "ours",
        "config": {"model": "fictitious-domain Poisson on ellipse, diagonal-PCG, delta=1e-6", "M": M, "N": N,
                   "global_batch": 1, "seq_len": nodes, "parallelism": f"{solver.grid[0]}x{solver.grid[1]} spatial "
                   "blocks, peer-memory halo + in-kernel reductions", "l2": "512 MiB flush buffer written between "
                   "timed solves; each solve re-assembles all arrays", "method": "persistent cooperative kernel",
                   "div": a.div, "ap_mode": a.ap_mode, "ap_recomputed": recomputes, "ctas_per_rank": info["ctas_per_rank"]},
        "samples": dev_s, "iterations": iters[-1], "expected_iterations": EXPECTED_ITERS.get((M, N)), "status": res["status_name"],
        "us_per_iter": us_iter, "l2_error_in_D": res["l2_in_D"], "max_error": res["max_abs"],
        # DRAM bytes per node and iteration on constant-coefficient tiles: r, p, p' | w, r, p', w', r' (64) when
        # phase C recomputes Ap, 16 more (Ap written and read) when it is stored
        "hbm_gbs_moved_per_gpu": bytes_node * nodes / (us_iter * 1e-6) / 1e9 / world, "bytes_per_node_iter": bytes_node,
        "wall_s_timed_region": t_wall,
        "e2e": {"value": e2e, "unit": "s", "samples": e2e_s, "h2d_bytes_per_step": info["h2d_bytes"], "d2h_bytes_per_step": info["d2h_bytes"],
                "what": "poisson_b200.solve(M, N): options pinned->device, assembly, PCG, error report, result "
                        "block device->pinned; wall clock, max over ranks"},
        "gpu_launches": launches, "clocks": clocks, "roofline": roof,
        "timers_s": {k: res[k] for k in ("t_assemble", "t_iterations", "t_gpu", "t_comm", "t_dot", "t_phase_ab", "t_phase_c",
                                           "t_red1", "t_red2")},
    }
    pb.clear_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0 and a.cli_cold:
        # after the timed regions and after the process group is gone: the other ranks return and exit now, so that the
        # CLI does not have to create its contexts and peer mappings next to N busy processes
        del flush
        torch.cuda.empty_cache()
        if world > 1:
            time.sleep(3.0)
        out["cli_cold"] = cli_cold_seconds(M, N, world)
    return out if rank == 0 else None


# --------------------------------------------------------------------------------------------------
_REF_BUCKETS = {
    "t_gpu": r"GPU compute time.*~ ([0-9.eE+-]+) s", "t_copy": r"Host<->Device copy time.*~ ([0-9.eE+-]+) s",
    "t_comm": r"MPI halo exchange time.*~ ([0-9.eE+-]+) s", "t_precond": r"Preconditioner CPU part time.*~ ([0-9.eE+-]+) s",
    "t_dot": r"Dot products time.*~ ([0-9.eE+-]+) s"}


def _grab(pat: str, txt: str) -> float:
    m = re.search(pat, txt)
    return float(m.group(1)) if m else float("nan")


def parse_reference_report(txt: str) -> dict:
    """Numbers of one stage-4 report (rank 0's stdout; format: S4:900-902, S4:969-978, S4:1026-1033)."""
    out = {"solver_s": _grab(r"Solver time \(MPI\+CUDA\)\s+~ ([0-9.eE+-]+)", txt),
           "total_s": _grab(r"Total Time=([0-9.eE+-]+) s", txt),
           "init_s": _grab(r"Init time \(program\)\s+~ ([0-9.eE+-]+)", txt)}
    m = re.search(r"Iter=(\d+)", txt)
    out["iters"] = int(m.group(1)) if m else -1
    out["converged"] = "Converged after" in txt
    out.update({k: _grab(pat, txt) for k, pat in _REF_BUCKETS.items()})
    return out


def reference_copy_bytes(M: int, N: int, world: int, iters: int) -> tuple[int, int]:
    """Host<->device bytes of one reference solve on its busiest rank, counted from its source: H2D = the three
    coefficient arrays (S4:752-757) + per iteration the halo rows it copies back in (S4:435-496); D2H = per
    iteration three 256 KiB partial-sum buffers (S4:779-792, two dots + reduce_diff) and the halo rows it
    stages out (S4:362-399), plus the initial dot (S4:837)."""
    px = int(world ** 0.5)
    while world % px:
        px -= 1
    py = world // px  # choose_process_grid, S4:82-86
    nx, ny = -(-(M - 1) // px), -(-(N - 1) // py)  # the largest block
    arrays = 3 * (nx + 2) * (ny + 2) * 8
    halo = 0
    if px > 1:
        halo += min(2, px - 1) * (ny + 2) * 8
    if py > 1:
        halo += min(2, py - 1) * (nx + 2) * 8
    part = 32768 * 8
    return arrays + iters * halo, part + iters * (3 * part + halo)


def reference_line(reports: list, walls: list, M: int, N: int, world: int, steps: int, warmup: int, clocks) -> dict:
    """The JSON line of the reference arm from the parsed reports of the timed runs (max over ranks is already in
    the program's own numbers: its timers are MPI_Reduce(MAX)'ed and its phase times sit between MPI_Barriers)."""
    solver_s = [r["solver_s"] for r in reports]
    total_s = [r["total_s"] for r in reports]
    last = reports[-1]
    iters = last["iters"]
    value = sum(solver_s) / len(solver_s)
    loop_s = last["t_gpu"] + last["t_copy"] + last["t_comm"] + last["t_dot"]
    h2d, d2h = reference_copy_bytes(M, N, world, max(iters, 0))
    return {
        "metric": f"pcg_time_to_converge_{M}x{N}", "value": value, "unit": "s", "n_gpus": world, "steps": steps,
        "warmup": warmup, "ms_per_step": 1e3 * value, "higher_is_better": False, "scaling": "strong",
        "vs_baseline": None, "dtype": "fp64", "data": "synthetic (analytic problem)", "impl": "reference",
        "config": {"model": "reference stage4 poisson_mpi_cuda2.cu, unmodified, nvcc -O3 -std=c++11 -arch=sm_100",
                   "M": M, "N": N, "global_batch": 1, "seq_len": (M - 1) * (N - 1),
                   "parallelism": f"{world} shim-MPI ranks, 1 GPU each (CUDA_VISIBLE_DEVICES)",
                   "l2": "every step is a new process: nothing is cached between steps",
                   "timer": "the program's own 'Solver time (MPI+CUDA)' (MPI_Wtime between barriers, S4:1009-1016; "
                            "includes its CPU assembly, cudaMalloc, H2D; CUDA context created in MPI_Init)"},
        "samples": solver_s, "iterations": iters, "expected_iterations": EXPECTED_ITERS.get((M, N)),
        "status": "converged" if last["converged"] else "not converged",
        "us_per_iter": 1e6 * loop_s / max(1, iters),
        "timers_s": {k: last[k] for k in _REF_BUCKETS},
        # End to end through the reference's only public interface, `prog M N`: its own 'Total Time' line
        # (S4:1026-1030: from after MPI_Init to after the solver's closing barrier).  The wall clock of the whole
        # process (start-up, CUDA context creation in MPI_Init, exit) is reported next to it.
        "e2e": {"value": sum(total_s) / len(total_s), "unit": "s", "samples": total_s,
                "process_wall_s": sum(walls) / len(walls), "process_wall_samples": walls,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "what": "the program's own 'Total Time' line (max over ranks by its barriers); bytes counted from "
                        "the source for the rank with the largest block"},
        "clocks": clocks, "gpu_launches": 0,
    }


def run_reference(a) -> dict:
    """The unmodified stage-4 program of the reference (baseline/_ref), one shim rank per torchrun rank."""
    rank, world, local = _env()
    exe = ROOT / "baseline" / "_ref" / "bin" / "poisson_mpi_cuda"
    if not exe.exists():
        try:
            sys.path.insert(0, str(ROOT / "baseline"))
            import build_ref

            if rank == 0:
                build_ref.build(["poisson_mpi_cuda"])
        except Exception as e:  # noqa: BLE001
            return {"impl": "reference", "unavailable": f"cannot build baseline/_ref: {e}"[:200]} if rank == 0 else None
        for _ in range(600):
            if exe.exists():
                break
            time.sleep(0.5)
    if not exe.exists():
        return {"impl": "reference", "unavailable": "baseline/_ref/bin/poisson_mpi_cuda missing"} if rank == 0 else None
    M, N = a.M, a.N
    port = os.environ.get("MASTER_PORT", "0")
    sampler = ClockSampler(local) if rank == 0 else None
    runs, walls = [], []
    for step in range(a.warmup + a.steps):
        if sampler and step == a.warmup:
            sampler.start()
        env = dict(os.environ, PB_MPI_RANK=str(rank), PB_MPI_SIZE=str(world), PB_MPI_JOB=f"bench_{port}_{step}",
                   CUDA_VISIBLE_DEVICES=str(local))
        t0 = time.perf_counter()
        p = subprocess.run([str(exe), str(M), str(N)], env=env, capture_output=True, text=True)
        wall = time.perf_counter() - t0
        if p.returncode != 0:
            return {"impl": "reference", "unavailable": f"reference exited {p.returncode}: {p.stderr[-160:]}"} if rank == 0 else None
        if rank == 0 and step >= a.warmup:
            runs.append(p.stdout)
            walls.append(wall)
    if rank != 0:
        return None
    clocks = sampler.finish() if sampler else None
    try:
        return reference_line([parse_reference_report(t) for t in runs], walls, M, N, world, a.steps, a.warmup, clocks)
    except Exception as e:  # noqa: BLE001  -- the arm must always print a line
        return {"impl": "reference", "unavailable": f"could not assemble the result line: {e!r}"[:200]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--M", type=int, default=M_DEFAULT)
    ap.add_argument("--N", type=int, default=N_DEFAULT)
    ap.add_argument("--grid", default=None, help="process grid: PxxPy, 'reference' (the reference's near-square\n                    choice) or default 'auto' (N x 1 slabs, see parallel/decomp.py)")
    ap.add_argument("--div", default="markstein", choices=["markstein", "true", "recip"],
                    help="stencil divisions by h1/h2: markstein = correctly rounded 3-instruction sequence (same "
                         "bits as 'true'), true = IEEE division, recip = multiply by the reciprocal")
    ap.add_argument("--tile-rows", type=int, default=0)
    ap.add_argument("--no-cli-cold", dest="cli_cold", action="store_false",
                    help="skip the extra cold run of the CLI (reported as `cli_cold`, outside every timed region)")
    ap.add_argument("--ap-mode", type=int, default=0, choices=[0, 1, 2],
                    help="constant-coefficient tiles: 1 store and reload Ap, 2 recompute it in phase C, 0 library default")
    a = ap.parse_args()
    # The contract is ONE JSON line on stdout.  Libraries (NCCL with NCCL_DEBUG=VERSION, the reference program's
    # launcher) write to file descriptor 1 behind Python's back, so everything but the final line goes to stderr.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    out = run_reference(a) if a.impl == "reference" else run_ours(a)
    sys.stdout.flush()
    if out is not None:
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    os.close(json_fd)


if __name__ == "__main__":
    main()
