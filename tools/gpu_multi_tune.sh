#!/bin/bash
# N-GPU overlap tuning: bench.py under different all-reduce settings.
N=${1:-2}
mkdir -p gpurun_out
L=gpurun_out/multi_tune_${N}.log
: > $L
P=29600
run() { P=$((P+1)); echo "=== $*" >> $L; env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --steps 30 --warmup 5 2>&1 | grep -E "^\{|rror" | python -c "
import sys, json
for l in sys.stdin:
    try:
        d = json.loads(l); print('ms_per_step %.3f  img/s %.1f  e2e %.1f  blocks %s' % (d['ms_per_step'], d['value'], d['e2e']['value'], d.get('allreduce',{}).get('comm_blocks')))
    except Exception: print(l.strip()[:300])
" >> $L; }
run B200DP_AR_ALGO=nvls B200DP_AR_BLOCKS=64
run B200DP_AR_ALGO=nvls B200DP_AR_BLOCKS=48
run B200DP_AR_ALGO=nvls B200DP_AR_BLOCKS=32
run B200DP_AR_ALGO=nvls B200DP_AR_BLOCKS=96
echo "=== sweep" >> $L
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29700 tools/allreduce_sweep.py 2>&1 | grep -E "^\{" | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print('%11d B  ours %s  nccl %.3f ms  busbw ours %.0f nccl %.0f GB/s' % (d['bytes'], {k[5:-3]: round(v, 3) for k, v in d.items() if k.startswith('ours_') and k.endswith('_ms')}, d['nccl_ms'], d['ours_busbw_gbs'], d['nccl_busbw_gbs']))" >> $L
cat $L
