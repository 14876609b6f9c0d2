#!/bin/bash
# time_variants.sh NAME...  — single-GPU bench (per-kernel table) with build/variants/lib_NAME.so ("base" = the in-tree library)
mkdir -p gpurun_out
for NAME in "$@"; do
  if [ "$NAME" = base ]; then unset VCR_GAUS_LIB; else export VCR_GAUS_LIB=$PWD/vcr-gaus_b200/build/variants/lib_$NAME.so; fi
  timeout 300 python bench.py --no-cpu-baseline --no-full-step --steps 40 > gpurun_out/var_$NAME.json 2> gpurun_out/var_$NAME.err || tail -5 gpurun_out/var_$NAME.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/var_$NAME.json').read().strip().splitlines()[-1])
k={r['kernel']:r['avg_launch_ms']*1e3 for r in d.get('kernels',[])}
print('%-10s ms/step %.4f  e2e %.4f  | bwd %.1f fwd %.1f radix %.1f pre_bwd %.1f pre %.1f dup %.1f us' % ('$NAME', d['ms_per_step'], d['e2e']['ms_per_step'], k.get('render_backward_kernel',0), k.get('render_forward_kernel',0), k.get('radix_pass_kernel',0), k.get('preprocess_backward_kernel',0), k.get('preprocess_kernel',0), k.get('duplicate_kernel',0)))
PY
done
