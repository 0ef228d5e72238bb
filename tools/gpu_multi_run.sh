#!/bin/bash
# Multi-GPU run (gpurun --gpus N): native collectives tests, bench at N (both arms).
N=${1:-2}
mkdir -p gpurun_out
L=gpurun_out/multi_run_${N}.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-300} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
[ -z "$SKIP_TESTS" ] && run python -m pytest -q -s -p no:cacheprovider -m gpu tests/test_gpu_multi.py
run $TR --nproc-per-node $N --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 $BENCH_ARGS
[ -n "$REF" ] && TMO=400 run $TR --nproc-per-node $N --master-port 29531 bench.py --impl reference --gpus $N --steps 20 --warmup 5
grep -E "passed|failed|rc=|rror|\[multi\] DDP" $L | cut -c1-300 | tail -20
grep -E "^\{" $L | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if d.get('impl') == 'reference':
        print('REF N=%d ms_per_step %.3f img/s %.2f e2e %.2f' % (d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'])); continue
    print('N=%d ms_per_step %.3f  img/s %.1f  e2e %.1f  sustained %.1f  bf16 %.3f ms  %s  identical=%s' % (d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d.get('sustained',{}).get('value',0), d.get('bf16_mode',{}).get('ms_per_step',0), {k: d.get('allreduce',{}).get(k) for k in ('fc_bucket_algo','comm_blocks','sm_partition')}, d.get('replicas_bit_identical')))"
