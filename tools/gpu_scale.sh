#!/usr/bin/env bash
# Multi-GPU call on an N-GPU box: 2-GPU exchange test, then bench at every N in "$@" for metric / C / D (both arms where cheap)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multimem_gpu.py -q -rs > gpurun_out/scale_mm_test.log 2>&1; echo "[mm_test] rc=$? $(tail -1 gpurun_out/scale_mm_test.log)"
for N in "$@"; do
  L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
  [ "$N" = 1 ] && L="python"
  for CFG in metric C D; do
    EXTRA="--steps 30 --warmup 10"; [ $CFG = C ] && EXTRA="--steps 300 --warmup 20"; [ $CFG = D ] && EXTRA="--steps 12 --warmup 4"
    timeout 900 $L bench.py --gpus $N --config $CFG $EXTRA --no-cpu-baseline --no-full-step > gpurun_out/scale_${CFG}_n$N.json 2> gpurun_out/scale_${CFG}_n$N.err
    echo "[$CFG N=$N] rc=$? $(python -c "
import json
try:
    d=json.loads(open('gpurun_out/scale_${CFG}_n$N.json').read().strip().splitlines()[-1]); print('%.3f ms/step %.1f views/s e2e %.3f' % (d['ms_per_step'], d['value'], d['e2e']['ms_per_step']), (d.get('exchange') or {}).get('chosen'), (d.get('exchange') or {}).get('verified'), d.get('gaussians_after_each_densification'))
except Exception as e: print('no json', e)")"
  done
done
if [ -n "$REF_N" ]; then
  for N in $REF_N; do
    L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518"
    timeout 900 $L bench.py --gpus $N --impl reference --steps 30 --warmup 10 > gpurun_out/scale_refmetric_n$N.json 2> gpurun_out/scale_refmetric_n$N.err
    echo "[ref metric N=$N] rc=$? $(head -c 300 gpurun_out/scale_refmetric_n$N.json)"
  done
fi
