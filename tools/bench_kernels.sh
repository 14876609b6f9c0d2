# full GPU test suite + single-GPU bench with the per-kernel table (prints a compact summary)
python -m pytest tests -x -q -m gpu 2>&1 | tail -6
VCR_BENCH_NO_FULL_STEP=1 python bench.py --no-cpu-baseline > gpurun_out/var_${1:-k}.json 2> gpurun_out/var_${1:-k}.err || tail -5 gpurun_out/var_${1:-k}.err
python - <<PY
import json
d=json.loads(open("gpurun_out/var_${1:-k}.json").read().strip().splitlines()[-1])
print("ms/step %.4f e2e %.4f launches %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["gpu_launches"]))
for r in d["kernels"]: print("%-32s %5.1f x %8.1f us = %.3f ms" % (r["kernel"], r["launches_per_step"], r["avg_launch_ms"]*1e3, r["ms_per_step"]))
PY
