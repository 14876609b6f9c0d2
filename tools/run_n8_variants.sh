# diagnosis runs at N = $1 (default 8): NCCL unchunked, multimem unchunked, exchange microbenchmark
N=${1:-8}
run() { # name, env...
  name=$1; shift
  env "$@" VCR_BENCH_DIAG=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --no-cpu-baseline --no-profile > gpurun_out/n${N}_$name.json 2> gpurun_out/n${N}_$name.err
  echo "== $name rc=$?"; grep DIAG gpurun_out/n${N}_$name.err | cut -c1-1500; python -c "
import json;d=json.loads(open('gpurun_out/n${N}_$name.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['value'],d['e2e']['ms_per_step'],d['clocks'])"
}
run nccl_c0 VCR_BENCH_CHUNKS=0 VCR_BENCH_TRANSPORT=nccl
run mm_c1 VCR_BENCH_CHUNKS=1
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 tools/mm_bench.py 2>&1 | grep -a "MMBENCH\|Error\|error" | head
