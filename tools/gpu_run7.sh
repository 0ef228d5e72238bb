#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/run7.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-300} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
run python -m pytest -q -p no:cacheprovider -m gpu tests/test_gpu_kernels.py -k "conv2 and cpasync"
run python tools/gpu_first_light.py tc3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29801 tools/ddp_breakdown.py 2>&1 | grep -E "^---|^  " >> $L
bash tools/gpu_multi_tune.sh 2 > /dev/null 2>&1; cat gpurun_out/multi_tune_2.log >> $L
grep -E "passed|failed|rc=|rror" $L | head; grep -vE "c_void_p|c_int" gpurun_out/first_light.txt | tail -18; grep -A16 "^---" $L | head -18; grep -E "^===|ms_per_step" gpurun_out/multi_tune_2.log
