#!/usr/bin/env bash
# One gpurun call: GPU test suite, then the named bench runs; every step under its own timeout, logs into gpurun_out/<tag>_*.
#   tools/gpu_batch.sh TAG [steps...]      steps: tests ubench metric ref A B refA refB
TAG=${1:-run}; shift
OUT=gpurun_out
mkdir -p $OUT
run() { # name timeout cmd...
  local name=$1 to=$2; shift 2
  local t0=$(date +%s)
  timeout $to "$@" > $OUT/${TAG}_$name.out 2> $OUT/${TAG}_$name.err
  local rc=$?
  echo "[$name] rc=$rc $(( $(date +%s) - t0 ))s" | tee -a $OUT/${TAG}_summary.txt
  return $rc
}
for step in "$@"; do
  case $step in
    tests)  run tests 1500 python -m pytest tests -m gpu -x -q; tail -5 $OUT/${TAG}_tests.out ;;
    testsall) run tests 1800 python -m pytest tests -m gpu -q; tail -15 $OUT/${TAG}_tests.out ;;
    ubench) nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o $OUT/ubench_f32x2 tools/ubench_f32x2.cu && run ubench 60 $OUT/ubench_f32x2; cat $OUT/${TAG}_ubench.out ;;
    metric) run metric 600 python bench.py --steps 30 --warmup 10; head -c 600 $OUT/${TAG}_metric.out; echo ;;
    metric1) VCR_RENDER_VERSION=1 run metric1 600 python bench.py --steps 30 --warmup 10 --no-cpu-baseline --no-full-step; head -c 300 $OUT/${TAG}_metric1.out; echo ;;
    metricq) run metricq 600 python bench.py --steps 30 --warmup 10 --no-cpu-baseline --no-full-step; head -c 300 $OUT/${TAG}_metricq.out; echo ;;
    ncu_render) # ncu --set full of one steady-state forward + backward compositing launch (after the same command ran clean)
        export VCR_BENCH_WARMUP_FLOOR=20
        CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-full-step --no-profile"
        $CMD > $OUT/${TAG}_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:render_ -s 30 -c 2 -f -o $OUT/${TAG}_render $CMD > $OUT/${TAG}_ncu.log 2>&1
        echo "[ncu_render] rc=$?" | tee -a $OUT/${TAG}_summary.txt; tail -3 $OUT/${TAG}_ncu.log
        unset VCR_BENCH_WARMUP_FLOOR ;;
    ncu_small) # ncu --set full of the HBM-side kernels in steady state (after the same command ran clean)
        export VCR_BENCH_WARMUP_FLOOR=20
        CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-full-step --no-profile"
        $CMD > $OUT/${TAG}_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k "regex:radix_pass|duplicate_sorted|preprocess|view_loss_backward|tile_order|scan_block" -s 60 -c 14 -f -o $OUT/${TAG}_small $CMD > $OUT/${TAG}_ncu.log 2>&1
        echo "[ncu_small] rc=$?" | tee -a $OUT/${TAG}_summary.txt; tail -3 $OUT/${TAG}_ncu.log
        unset VCR_BENCH_WARMUP_FLOOR ;;
    launches) # per-launch durations of two steady-state steps (after the same command ran clean)
        CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-full-step --no-profile"
        $CMD > $OUT/${TAG}_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launch_list.csv $CMD > $OUT/${TAG}_ncu_list.log 2>&1
        echo "[launches] rc=$?" | tee -a $OUT/${TAG}_summary.txt ;;
    ref)    run ref 600 python bench.py --impl reference --steps 30 --warmup 10; head -c 600 $OUT/${TAG}_ref.out; echo ;;
    A)      run A 300 python bench.py --config A --steps 50 --warmup 10; head -c 400 $OUT/${TAG}_A.out; echo ;;
    refA)   run refA 300 python bench.py --config A --impl reference --steps 50 --warmup 10; head -c 400 $OUT/${TAG}_refA.out; echo ;;
    B)      run B 600 python bench.py --config B --steps 49 --warmup 10 --no-cpu-baseline; head -c 400 $OUT/${TAG}_B.out; echo ;;
    refB)   run refB 600 python bench.py --config B --impl reference --steps 49 --warmup 10; head -c 400 $OUT/${TAG}_refB.out; echo ;;
    C)      run C 900 python bench.py --config C --steps 300 --warmup 20 --no-cpu-baseline; head -c 600 $OUT/${TAG}_C.out; echo ;;
    refC)   run refC 1200 python bench.py --config C --impl reference --steps 300 --warmup 20; head -c 600 $OUT/${TAG}_refC.out; echo ;;
    *) echo "unknown step $step" ;;
  esac
done
tail -3 $OUT/${TAG}_*.err 2>/dev/null | tail -40
