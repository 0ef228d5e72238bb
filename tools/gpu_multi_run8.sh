#!/bin/bash
# 8-GPU confirmation: multi tests, bench (default partition, 8, 12), N=4, per-kernel breakdown
N=${1:-8}
mkdir -p gpurun_out
L=gpurun_out/multi_run8_${N}.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-240} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run python -m pytest -q -s -p no:cacheprovider -m gpu tests/test_gpu_multi.py
run $TR --nproc-per-node $N --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5
run env B200DP_RESERVED_SMS=8 $TR --nproc-per-node $N --master-port 29521 bench.py --gpus $N --steps 30 --warmup 5
run env B200DP_RESERVED_SMS=12 $TR --nproc-per-node $N --master-port 29522 bench.py --gpus $N --steps 30 --warmup 5
run env B200DP_RESERVED_SMS=4 $TR --nproc-per-node $N --master-port 29523 bench.py --gpus $N --steps 30 --warmup 5
H=$((N/2))
run $TR --nproc-per-node $H --master-port 29514 bench.py --gpus $H --steps 30 --warmup 5
run $TR --nproc-per-node $N --master-port 29531 tools/ddp_breakdown.py
run env B200DP_DEBUG_SKIP_FC_ALLREDUCE=1 $TR --nproc-per-node $N --master-port 29532 tools/ddp_breakdown.py
grep -E "passed|failed|rc=|rror|\[multi\] DDP" $L | cut -c1-300 | tail -20
grep -E "^\{\"metric" $L | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('N=%d ms_per_step %.3f  img/s %.1f  e2e %.1f  %s' % (d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], {k: d.get('allreduce',{}).get(k) for k in ('fc_bucket_algo','comm_blocks','sm_partition')}))"
grep -E "^---|^  " $L | head -50
