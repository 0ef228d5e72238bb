#!/bin/bash
# Multi-GPU run (gpurun --gpus N): native collectives tests, bench at N (and N/2), all-reduce sweep, reference arm.
N=${1:-2}
mkdir -p gpurun_out
L=gpurun_out/multi_run_${N}.log
: > $L
nvidia-smi topo -m >> $L 2>&1
run() { echo "=== $*" >> $L; timeout ${TMO:-300} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run python -m pytest -q -s -p no:cacheprovider -m gpu tests/test_gpu_multi.py
run $TR --nproc-per-node $N --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5
B200DP_AR_BLOCKS=64 run env B200DP_AR_BLOCKS=64 $TR --nproc-per-node $N --master-port 29521 bench.py --gpus $N --steps 30 --warmup 5
if [ $N -ge 4 ]; then
  H=$((N/2))
  run $TR --nproc-per-node $H --master-port 29514 bench.py --gpus $H --steps 30 --warmup 5
fi
run $TR --nproc-per-node $N --master-port 29512 tools/allreduce_sweep.py
TMO=400 run $TR --nproc-per-node $N --master-port 29513 bench.py --impl reference --gpus $N --steps 5 --warmup 3 --stock-loader-steps 0
cp gpurun_out/allreduce_sweep_${N}gpu.jsonl gpurun_out/ 2>/dev/null
grep -E "passed|failed|rc=|rror|^\{\"metric|^\{\"impl|\[multi\] DDP" $L | cut -c1-420 | tail -40
