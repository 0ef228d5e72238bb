#!/bin/bash
# ncu capture of one kernel (plain run first, same command line, as the recipe requires)
K=$1; REGEX=$2; VER=${3:-3}; SKIP=${4:-2}
mkdir -p gpurun_out
python tools/run_one_kernel.py $K $VER 3 > gpurun_out/plain_$K.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$REGEX -s $SKIP -c 1 -f -o gpurun_out/prof_${K}_v$VER \
    python tools/run_one_kernel.py $K $VER 3 > gpurun_out/ncu_$K.log 2>&1
echo "rc=$?"; cat gpurun_out/plain_$K.log | tail -2; tail -3 gpurun_out/ncu_$K.log; ls -la gpurun_out/*.ncu-rep
