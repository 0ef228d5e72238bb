#!/bin/bash
# default (vt) tf32 evidence: whole-step ncu table, precision error table (both conv1 forward kernels), loss curves
mkdir -p gpurun_out
python tools/ncu_step.py tf32 > gpurun_out/step_plain_tf32.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -f -o /tmp/prof_step_tf32 \
    python tools/ncu_step.py tf32 > gpurun_out/step_ncu_tf32.log 2>&1
python tools/ncu_step_summary.py /tmp/prof_step_tf32.ncu-rep > gpurun_out/step_ncu_tf32_vt.txt 2> gpurun_out/step_ncu_tf32_err.log
tail -1 gpurun_out/step_plain_tf32.log; tail -1 gpurun_out/step_ncu_tf32_vt.txt
timeout 300 python tools/dev/tf32_err.py > gpurun_out/precision_error_vt.txt 2>&1; tail -2 gpurun_out/precision_error_vt.txt
timeout 600 python tools/convergence.py 100 3000 1e-6 > gpurun_out/conv_1e-6_vt.log 2>&1; tail -2 gpurun_out/conv_1e-6_vt.log
cp gpurun_out/convergence_3000_lr1e-06.txt gpurun_out/convergence_3000_lr1e-06_vt.txt
python bench.py --steps 20 --warmup 5 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; cut -c1-400 gpurun_out/bench1.json
timeout 300 python -m pytest tests/test_gpu_engine.py tests/test_gpu_convergence.py -q -x 2>&1 | tail -2
