#!/bin/bash
# A/B of DDP knobs at N GPUs: gpu_multi_ab.sh N "ENV=.. ENV2=.." "ENV=.." ...   (bench, one precision, short)
N=$1; shift
mkdir -p gpurun_out
L=gpurun_out/multi_ab_${N}.log
: > $L
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
P=29600
if [ -n "$TESTS" ]; then echo "=== tests" >> $L; timeout 600 python -m pytest -q -s -p no:cacheprovider -m gpu tests/test_gpu_multi.py >> $L 2>&1; echo "=== rc=$?" >> $L; fi
for cfg in "$@"; do
  P=$((P+1))
  echo "=== cfg: $cfg" >> $L
  timeout 200 env $cfg $TR --nproc-per-node $N --master-port $P bench.py --gpus $N --steps 20 --warmup 5 --precision ${PREC:-tf32} --sustained-seconds 0 >> $L 2>&1
  echo "=== rc=$?" >> $L
done
grep -E "passed|failed|=== cfg|rc=|rror|timed out" $L | cut -c1-200 | tail -40
grep -E "^=== cfg|^\{" $L | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('=== cfg'): print(l.strip()); continue
    d = json.loads(l)
    print('   N=%d ms_per_step %.3f  img/s %.1f  e2e %.1f  %s identical=%s' % (d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d.get('allreduce',{}).get('fc_bucket_algo'), d.get('replicas_bit_identical')))"
