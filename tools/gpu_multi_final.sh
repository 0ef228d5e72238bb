#!/bin/bash
# Final multi-GPU confirmation with default knobs: multi tests, bench at N, N/2, N/4
N=${1:-8}
mkdir -p gpurun_out
L=gpurun_out/multi_final_${N}.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-240} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run python -m pytest -q -s -p no:cacheprovider -m gpu tests/test_gpu_multi.py
P=29540
for n in $N $((N/2)) $((N/4)); do
  [ $n -ge 2 ] || continue
  P=$((P+1))
  run $TR --nproc-per-node $n --master-port $P bench.py --gpus $n --steps 40 --warmup 5
done
grep -E "passed|failed|rc=|rror|\[multi\] DDP" $L | cut -c1-300 | tail -12
grep -E "^\{\"metric" $L | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('N=%d ms_per_step %.3f  img/s %.1f  e2e %.1f  %s clocks %s' % (d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], {k: d.get('allreduce',{}).get(k) for k in ('fc_bucket_algo','comm_blocks','sm_partition')}, d.get('clocks')))"
