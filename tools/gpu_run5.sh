#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/run5.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-300} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
PT="python -m pytest -q -s -p no:cacheprovider -m gpu"
run $PT tests/test_gpu_kernels.py -k "conv2_fwd_tc and cpasync"
run $PT tests/test_gpu_kernels.py -k "conv2_dgrad_tc and cpasync"
run $PT tests/test_gpu_kernels.py -k "conv2_wgrad_tc and cpasync"
run python tools/gpu_first_light.py tc3
cp gpurun_out/first_light.txt gpurun_out/first_light_v3.txt
grep -E "passed|failed|rc=|rror" $L | tail -30
grep -vE "c_void_p|c_int" gpurun_out/first_light_v3.txt
