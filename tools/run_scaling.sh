# bench at N = $1 with diagnostics; prints the JSON essentials
N=${1:-2}; shift
name=${1:-default}; shift
if [ "$N" = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"; fi
env "$@" VCR_BENCH_DIAG=1 timeout 400 $L bench.py --gpus $N --no-cpu-baseline --no-profile > gpurun_out/n${N}_$name.json 2> gpurun_out/n${N}_$name.err
echo "== N=$N $name rc=$?"; grep DIAG gpurun_out/n${N}_$name.err | cut -c1-1500; python -c "
import json;d=json.loads(open('gpurun_out/n${N}_$name.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['value'],d['e2e']['ms_per_step'],d['clocks'],d.get('exchange'))"
