#!/bin/bash
# Measurement campaign on N GPUs of one box: bench/tools/campaign.sh N [steps] -> gpurun_out/campaign_N/*.
# Ours and the reference (both arms of bench.py) at 4096 x 4096 and 800 x 1200, the CLI from a cold start, and -- at
# N >= 2 -- the communication sweeps (ours / NCCL and the reference's own paths) and the multi-GPU tests; at N = 8 the
# weak-scaling configurations.
N=${1:-1}; STEPS=${2:-3}
cd "$(dirname "$0")/../.."
OUT=gpurun_out/campaign_$N; mkdir -p $OUT
if [ "$N" = 1 ]; then TR="python"; else TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"; fi
nvidia-smi --query-gpu=index,name,clocks.max.sm,power.limit --format=csv,noheader > $OUT/gpus.txt
# the box's peer-to-peer matrix as CUDA reports it, and the process grid derived from it
python - $N > $OUT/p2p_topology.json 2>/dev/null <<'PY'
import json, sys
sys.path.insert(0, ".")
from poisson_ellipse_openmp_mpi_cuda_b200.parallel.topology import grid_from_topology, probe_p2p
n = int(sys.argv[1])
m = probe_p2p(range(n))
px, py, order, kind = grid_from_topology(m["perf"], 4096)
print(json.dumps({"n_gpus": n, "performance_rank": m["perf"], "access": m["access"], "native_atomics": m["atomics"],
                  "grid_for_4096": [px, py], "device_of_rank": order, "kind": kind}))
PY
# QUICK=1: our arm and the CLI only (a re-measurement after a kernel change)
if [ "$N" != 1 ] && [ -z "$QUICK" ]; then timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > $OUT/pytest_multi_gpu.log 2>&1; tail -2 $OUT/pytest_multi_gpu.log; fi
timeout 300 $TR bench.py --gpus $N --steps $STEPS --warmup 3 > $OUT/bench_ours_4096.json 2> $OUT/bench_ours_4096.err
timeout 300 $TR bench.py --gpus $N --steps $STEPS --warmup 3 --M 800 --N 1200 > $OUT/bench_ours_800x1200.json 2> $OUT/bench_ours_800.err
if [ -z "$QUICK" ]; then
timeout 600 $TR bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $OUT/bench_reference_4096.json 2> $OUT/bench_ref_4096.err
timeout 300 $TR bench.py --impl reference --gpus $N --steps 2 --warmup 1 --M 800 --N 1200 > $OUT/bench_reference_800x1200.json 2> $OUT/bench_ref_800.err
fi
B=poisson_ellipse_openmp_mpi_cuda_b200/bin/poisson_b200
for g in "4096 4096" "800 1200"; do set -- $g; ( time $B $1 $2 --gpus $N ) > $OUT/cli_${1}x${2}.txt 2>&1; done
if [ "$N" != 1 ] && [ -z "$QUICK" ]; then
  SWEEP_ITERS=500 timeout 300 $TR bench/halo_allreduce_sweep.py > $OUT/halo_allreduce_sweep.jsonl 2> $OUT/sweep.err
  timeout 300 python bench/ref_comm_sweep.py --ranks $N --iters 100 > $OUT/ref_comm_sweep.jsonl 2> $OUT/ref_comm.err
fi
if [ "$N" = 8 ] && [ -z "$QUICK" ]; then timeout 400 $TR bench/scaling.py --configs weak --reps 1 > $OUT/weak_scaling.jsonl 2> $OUT/weak.err; fi
for f in $OUT/bench_*.json; do python - $f <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], d.get("impl"), "value=%.4f" % d["value"], "us/it=%.2f" % d["us_per_iter"], "iters=%s" % d["iterations"],
          "e2e=%.4f" % d["e2e"]["value"], "roof=%s" % (d.get("roofline") or {}).get("fraction_of_roofline"), (d.get("clocks") or {}).get("sm_mhz"), (d.get("cli_cold") or ""))
except Exception as e:
    print(sys.argv[1], "UNREADABLE", e)
PY
done
grep -h "Total Time\|Solver time\|real" $OUT/cli_*.txt
tail -3 $OUT/weak_scaling.jsonl 2>/dev/null | cut -c1-400
