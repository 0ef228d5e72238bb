#!/bin/bash
# round 2 dev loop: fp32/tf32 kernels + engine tests, then per-kernel breakdowns
mkdir -p gpurun_out
L=gpurun_out/r2_first.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-600} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
run python tools/dev/dbg_wgrad_tf32.py
run python -m pytest tests/test_gpu_kernels_f32.py -q -p no:cacheprovider -s
run python -m pytest tests/test_gpu_engine.py -x -q -p no:cacheprovider -s
run python tools/step_breakdown.py --precision tf32
grep -E "passed|failed|rc=|rror|^---|^  [a-z_0-9]+ +[0-9.]+$|numerics f32\] conv1_wgrad|engine/tf32" $L | head -120
