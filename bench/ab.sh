#!/bin/bash
# A/B of native-library variants on the same box in the same call: bench/ab.sh "libA libB ..." [bench.py args...]
# (names under build/variants/, see bench/experiments/build_variant.sh; "." = the in-tree library).  Two rounds,
# interleaved; prints value, us/iteration and the phase / reduction clocks per variant.
LIBS=$1; shift
ROOT=$(cd "$(dirname "$0")/.." && pwd)
for round in 1 2; do
  for L in $LIBS; do
    if [ "$L" = "." ]; then unset POISSON_B200_LIB; else export POISSON_B200_LIB=$ROOT/build/variants/$L/libpoisson_b200.so; fi
    python $ROOT/bench.py "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); t=d['timers_s']; it=d['iterations']
print('%-18s' % '$L', 'value=%.4f us/it=%.2f  ab=%.2f c=%.2f r1=%.2f r2=%.2f us/it  iters=%d clocks=%s' % (d['value'], d['us_per_iter'], 1e6*t['t_phase_ab']/it, 1e6*t['t_phase_c']/it, 1e6*t['t_red1']/it, 1e6*t['t_red2']/it, it, (d.get('clocks') or {}).get('sm_mhz')))"
  done
done
