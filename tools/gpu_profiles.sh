#!/bin/bash
# Refresh the judged evidence under profiles/ on one GPU: whole-step ncu tables (after a plain run of the same program),
# single-step gradient error vs cuDNN-TF32, full-size loss curves vs the unmodified reference, the OOM experiment.
mkdir -p gpurun_out
for p in tf32 bf16; do
  python tools/ncu_step.py $p > gpurun_out/step_plain_$p.log 2>&1 && \
  ncu --set full --clock-control none --import-source on --profile-from-start off -f -o /tmp/prof_step_$p \
      python tools/ncu_step.py $p > gpurun_out/step_ncu_$p.log 2>&1
  python tools/ncu_step_summary.py /tmp/prof_step_$p.ncu-rep > gpurun_out/step_ncu_$p.txt 2> gpurun_out/step_ncu_${p}_err.log
  tail -1 gpurun_out/step_plain_$p.log; tail -2 gpurun_out/step_ncu_$p.txt
done
timeout 300 python tools/dev/tf32_err.py > gpurun_out/precision_error.txt 2>&1; tail -3 gpurun_out/precision_error.txt
timeout 600 python tools/convergence.py 100 3000 1e-6 > gpurun_out/conv_1e-6.log 2>&1; tail -3 gpurun_out/conv_1e-6.log
timeout 600 python tools/convergence.py 60 3000 1e-4 > gpurun_out/conv_1e-4.log 2>&1; tail -3 gpurun_out/conv_1e-4.log
timeout 600 python tools/oom_repro.py > gpurun_out/oom.log 2>&1; tail -5 gpurun_out/oom.log
