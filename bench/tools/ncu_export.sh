#!/bin/bash
# Text summaries of the captures in gpurun_out/ -> profiles/r2/ncu/ (the metrics B200_PROFILING.md names).
cd "$(dirname "$0")/../.."
mkdir -p profiles/r2/ncu
for rep in gpurun_out/ncu_*.ncu-rep; do
  base=$(basename $rep .ncu-rep)
  gzip -c $rep > profiles/r2/ncu/$base.ncu-rep.gz
  ncu -i $rep --page raw --csv 2>/dev/null > /tmp/$base.raw.csv
  python - /tmp/$base.raw.csv <<'PY' > profiles/r2/ncu/$base.summary.txt
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
if len(rows) < 3:
    print("no data"); sys.exit(0)
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed.sum",
        "smsp__issue_active.avg.pct", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__grid_size", "launch__block_size",
        "smsp__pcsamp_warps_issue_stalled_barrier", "smsp__pcsamp_warps_issue_stalled_long_scoreboard", "smsp__pcsamp_warps_issue_stalled_math_pipe_throttle",
        "smsp__pcsamp_warps_issue_stalled_membar", "smsp__pcsamp_warps_issue_stalled_short_scoreboard", "smsp__pcsamp_warps_issue_stalled_wait",
        "smsp__pcsamp_sample_buffer_total", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"]
for w in want:
    for i, h in enumerate(hdr):
        if h == w or h.endswith("." + w):
            print(f"{h} [{units[i]}] = {vals[i]}")
PY
done
ls -la profiles/r2/ncu
