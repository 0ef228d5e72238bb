#!/bin/bash
# ncu capture of one kernel (plain run first, same command line, as the recipe requires)
# usage: gpu_ncu.sh <kernel key of run_one_kernel.py> <kernel-name regex> [tc version] [launch skip] [precision]
K=$1; REGEX=$2; VER=${3:-5}; SKIP=${4:-2}; PREC=${5:-bf16}
mkdir -p gpurun_out
export B200DP_PRECISION=$PREC
python tools/run_one_kernel.py $K $VER 3 > gpurun_out/plain_${K}_$PREC.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$REGEX -s $SKIP -c 1 -f -o gpurun_out/prof_${K}_${PREC} \
    python tools/run_one_kernel.py $K $VER 3 > gpurun_out/ncu_${K}_$PREC.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/plain_${K}_$PREC.log; tail -3 gpurun_out/ncu_${K}_$PREC.log
