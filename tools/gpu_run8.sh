#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/run8.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-300} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
run python -m pytest -q -p no:cacheprovider -m gpu tests/test_gpu_kernels.py -k "conv1_fwd"
run python -m pytest -q -p no:cacheprovider -m gpu tests/test_gpu_engine.py
run python tools/gpu_first_light.py tc3
grep -E "passed|failed|rc=|rror" $L | head; grep -vE "c_void_p|c_int" gpurun_out/first_light.txt | tail -18
