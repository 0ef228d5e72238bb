#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
L=gpurun_out/multi_tune_${N}.log
: > $L
P=29600
run() { P=$((P+1)); echo "=== $*" >> $L; env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --steps 30 --warmup 5 2>&1 | grep -E "^\{|rror" | python -c "
import sys, json
for l in sys.stdin:
    try:
        d = json.loads(l); print('ms_per_step %.3f  img/s %.1f  e2e %.1f  blocks %s launches/step %s' % (d['ms_per_step'], d['value'], d['e2e']['value'], d.get('allreduce',{}).get('comm_blocks'), d.get('gpu_launches_per_step')))
    except Exception: print(l.strip()[:300])
" >> $L; }
echo "=== tests" >> $L
timeout 300 python -m pytest -q -s -p no:cacheprovider -m gpu tests/test_gpu_multi.py 2>&1 | grep -E "passed|failed|rror|DDP vs" >> $L
run B200DP_AR_BLOCKS=32
run B200DP_AR_BLOCKS=64
run B200DP_AR_BLOCKS=16
run B200DP_DEBUG_SKIP_FC_ALLREDUCE=1
cat $L
