#!/bin/bash
# compute-sanitizer over the persistent TMA kernel: racecheck (shared-memory hazards
# between the TMA pipeline, the in-place p update and the stencil), synccheck (barriers under divergence: the
# CTA-uniform abort path), memcheck.  Logs go to gpurun_out/sanitizer_*.log; copy them to profiles/sanitizer/.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
for tool in synccheck racecheck memcheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python bench/tools/sanitize_target.py > gpurun_out/sanitizer_$tool.log 2>&1
  echo "$tool rc=$?" | tee -a gpurun_out/sanitizer_$tool.log
  tail -4 gpurun_out/sanitizer_$tool.log
done
