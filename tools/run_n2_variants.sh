python -m pytest tests/test_multimem_gpu.py tests/test_view_parallel_gpu.py -x -q 2>&1 | tail -8
run() { # name, env...
  name=$1; shift
  env "$@" VCR_BENCH_DIAG=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --no-cpu-baseline --no-profile > gpurun_out/n2_$name.json 2> gpurun_out/n2_$name.err
  echo "== $name rc=$?"; grep DIAG gpurun_out/n2_$name.err | cut -c1-700; python -c "
import json;d=json.loads(open('gpurun_out/n2_$name.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['value'],d['e2e']['ms_per_step'])"
}
run mm_c4 VCR_BENCH_CHUNKS=4
run mm_c1 VCR_BENCH_CHUNKS=1
run mm_c8 VCR_BENCH_CHUNKS=8
run mm_c4_b148 VCR_BENCH_CHUNKS=4 VCR_MM_BLOCKS=148
run mm_c4_b592 VCR_BENCH_CHUNKS=4 VCR_MM_BLOCKS=592
tail -5 gpurun_out/n2_mm_c4.err
