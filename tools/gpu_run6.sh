#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/run6.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-300} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
PT="python -m pytest -q -s -p no:cacheprovider -m gpu"
run $PT tests/test_gpu_kernels.py -k "conv1_wgrad or conv1_fwd"
run $PT tests/test_gpu_engine.py
run python tools/gpu_first_light.py tc3
grep -E "passed|failed|rc=|rror|numerics.*conv1" $L | tail -30
grep -vE "c_void_p|c_int" gpurun_out/first_light.txt
