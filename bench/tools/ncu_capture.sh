#!/bin/bash
# One ncu capture of the persistent TMA kernel per configuration (B200_PROFILING.md: --set full, no clock control,
# source import), after the same command has run once without ncu.  Reports land in gpurun_out/; summaries are
# exported with bench/tools/ncu_export.sh and committed under profiles/.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
run() {  # name M N iterations [Px Py]
  name=$1; shift
  python bench/tools/ncu_target.py "$@" || return 1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:pcg_persistent_tma -c 1 -f \
    -o gpurun_out/$name python bench/tools/ncu_target.py "$@" > gpurun_out/$name.log 2>&1
  echo "$name rc=$?"; tail -2 gpurun_out/$name.log
}
# usage: ncu_capture.sh [suffix [only-first]]   (suffix goes into the report names, e.g. _rows)
SUF=$1
run ncu_tma_4096x4096_10it$SUF 4096 4096 10
[ -n "$2" ] && exit 0
run ncu_tma_4096x512_20it$SUF 4096 512 20
run ncu_tma_virtual2x1_4096x1024_10it$SUF 4096 1024 10 2 1
