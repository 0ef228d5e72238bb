#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/run10.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-600} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
run python -m pytest -q -p no:cacheprovider -m gpu tests/test_gpu_kernels.py
run python -m pytest -q -p no:cacheprovider -m gpu tests/test_gpu_engine.py
run python tools/gpu_first_light.py tc3
grep -E "passed|failed|rc=|rror" $L | head -20; grep -vE "c_void_p|c_int" gpurun_out/first_light.txt | tail -18
