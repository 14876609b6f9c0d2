#!/usr/bin/env bash
# Config E: rasterizer fwd+bwd sweep, ours and the reference build, one JSON line per point into gpurun_out/sweep_*.json
mkdir -p gpurun_out
export VCR_BENCH_WARMUP_FLOOR=20
for RES in "1920 1080" "3840 2160"; do
  set -- $RES; W=$1; H=$2
  for P in ${SWEEP_P:-100000 300000 1000000 3000000 10000000}; do
    for IMPL in ${SWEEP_IMPLS:-ours reference}; do
      f=gpurun_out/sweep_${IMPL}_${P}_${W}x${H}
      timeout 900 python bench.py --config E --impl $IMPL --gaussians $P --width $W --height $H --views 4 --steps 12 --warmup 4 --no-cpu-baseline --no-full-step > $f.json 2> $f.err
      echo "[$IMPL P=$P ${W}x${H}] rc=$? $(python -c "
import json,sys
try:
    d=json.loads(open('$f.json').read().strip().splitlines()[-1]); print('%.3f ms/view' % d['ms_per_step'], 'R=%s' % d.get('scene',{}).get('R'))
except Exception as e: print('no json', e)")"
    done
  done
done
