#!/bin/bash
# Development run on a B200 box: each test group in its own process (a trap in one kernel
# must not poison the CUDA context of the others), everything logged to gpurun_out/.
mkdir -p gpurun_out
L=gpurun_out/dev_run.log
: > $L
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv >> $L 2>&1
run() { echo "=== $*" >> $L; timeout 300 "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
PT="python -m pytest -q -s -p no:cacheprovider -m gpu"
run $PT tests/test_gpu_kernels.py -k "resize or conv1_fwd or bn_finalize or bn_relu_pool or fc_and_ce or pool_bn_bwd or conv1_wgrad or sgd"
run $PT tests/test_gpu_kernels.py -k "conv2_fwd_tc"
run $PT tests/test_gpu_kernels.py -k "conv2_dgrad_tc"
run $PT tests/test_gpu_kernels.py -k "conv2_wgrad_tc"
run $PT tests/test_gpu_engine.py -k "torch"
run $PT tests/test_gpu_engine.py -k "not torch"
run python tools/gpu_first_light.py torch
cp gpurun_out/first_light.txt gpurun_out/first_light_torch.txt 2>/dev/null
run python tools/gpu_first_light.py tc
grep -E "passed|failed|rc=|error" $L | tail -40
