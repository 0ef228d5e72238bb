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
        d = json.loads(l); print('ms_per_step %.3f  img/s %.1f  e2e %.1f  %s blocks %s' % (d['ms_per_step'], d['value'], d['e2e']['value'], d.get('allreduce',{}).get('fc_bucket_algo'), d.get('allreduce',{}).get('comm_blocks')))
    except Exception: print(l.strip()[:300])
" >> $L; }
echo "=== tests" >> $L
timeout 300 python -m pytest -q -s -p no:cacheprovider -m gpu tests/test_gpu_multi.py 2>&1 | grep -E "passed|failed|rror|DDP vs" >> $L
run B200DP_AR_BLOCKS=64
run B200DP_RESERVED_SMS=8
run B200DP_RESERVED_SMS=12
run B200DP_RESERVED_SMS=16
run B200DP_RESERVED_SMS=24
echo "=== single" >> $L
python bench.py --steps 30 --warmup 5 2>&1 | grep -E "^\{" | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('ms_per_step %.3f  img/s %.1f e2e %.1f' % (d['ms_per_step'], d['value'], d['e2e']['value']))" >> $L
python -m pytest -q -p no:cacheprovider -m gpu tests/test_gpu_kernels.py -k "conv1_fwd" 2>&1 | tail -1 >> $L
B200DP_RESERVED_SMS=16 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29811 tools/ddp_breakdown.py 2>&1 | grep -E "^---|^  " | head -8 >> $L
cat $L
