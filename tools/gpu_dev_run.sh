#!/bin/bash
# Development run on a B200 box: each test group in its own process (a trap in one kernel
# must not poison the CUDA context of the others), everything logged to gpurun_out/.
mkdir -p gpurun_out
L=gpurun_out/dev_run.log
: > $L
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv >> $L 2>&1
run() { echo "=== $*" >> $L; timeout ${TMO:-300} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
PT="python -m pytest -q -s -p no:cacheprovider -m gpu"
run $PT tests/test_gpu_kernels.py
run $PT tests/test_gpu_engine.py
run python -c "import __graft_entry__ as g; g.smoke()"
run python bench.py --steps 20 --warmup 5
TMO=600 run python bench.py --impl reference --steps 10 --warmup 3 --stock-loader-steps 3
run python tools/gpu_first_light.py tc
grep -E "passed|failed|rc=|error|^\{" $L | tail -40
