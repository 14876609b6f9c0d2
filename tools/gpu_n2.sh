#!/usr/bin/env bash
# Two-GPU call: the multicast-exchange test, then the bench at N=2 (exchange self-check in the JSON line).
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multimem_gpu.py -q -rs > gpurun_out/n${N}_mm_test.log 2>&1; echo "[mm_test] rc=$?"; tail -4 gpurun_out/n${N}_mm_test.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 10 --no-cpu-baseline > gpurun_out/n${N}_bench.json 2> gpurun_out/n${N}_bench.err
echo "[bench N=$N] rc=$?"; head -c 400 gpurun_out/n${N}_bench.json; echo; tail -3 gpurun_out/n${N}_bench.err
