#!/bin/bash
# Multi-GPU development run (gpurun --gpus N): native collectives, DDP engine, bench at N.
N=${1:-2}
mkdir -p gpurun_out
L=gpurun_out/multi_run_${N}.log
: > $L
nvidia-smi topo -m >> $L 2>&1
run() { echo "=== $*" >> $L; timeout ${TMO:-300} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run python -m pytest -q -s -p no:cacheprovider -m gpu tests/test_gpu_multi.py
run $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5
run $TR --master-port 29512 tools/allreduce_sweep.py
TMO=400 run $TR --master-port 29513 bench.py --impl reference --gpus $N --steps 5 --warmup 3 --stock-loader-steps 0
grep -E "passed|failed|rc=|rror|^\{|\[multi\]" $L | cut -c1-600 | tail -60
