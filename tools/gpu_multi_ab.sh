#!/bin/bash
# A/B of DDP knobs at N GPUs: usage gpu_multi_ab.sh N "ENV1=a ENV2=b" "ENV1=c" ...
N=$1; shift
mkdir -p gpurun_out
L=gpurun_out/multi_ab_${N}.log
: > $L
P=29700
for cfg in "$@"; do
  P=$((P+1))
  echo "=== $cfg" >> $L
  env $cfg timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --steps 40 --warmup 5 2>&1 | grep -E "^\{|rror" | python -c "
import sys, json
for l in sys.stdin:
    try:
        d = json.loads(l); print('ms_per_step %.3f  img/s %.1f  e2e %.1f  %s' % (d['ms_per_step'], d['value'], d['e2e']['value'], {k: d.get('allreduce',{}).get(k) for k in ('comm_blocks','sm_partition')}))
    except Exception: print(l.strip()[:300])
" >> $L
done
cat $L
