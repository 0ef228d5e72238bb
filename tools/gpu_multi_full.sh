#!/bin/bash
# Everything multi-GPU at N GPUs in one call: tests, bench both arms, all-reduce sweep, scaled-batch bench (bf16).
N=${1:-2}
mkdir -p gpurun_out
L=gpurun_out/multi_full_${N}.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-300} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
[ -z "$SKIP_TESTS" ] && TMO=500 run python -m pytest -q -s -p no:cacheprovider -m gpu tests/test_gpu_multi.py
run $TR --nproc-per-node $N --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5
[ -n "$REF" ] && TMO=400 run $TR --nproc-per-node $N --master-port 29531 bench.py --impl reference --gpus $N --steps 20 --warmup 5
[ -n "$SWEEP" ] && run $TR --nproc-per-node $N --master-port 29541 tools/allreduce_sweep.py
[ -n "$SCALED" ] && run $TR --nproc-per-node $N --master-port 29551 bench.py --gpus $N --batch $SCALED --precision bf16 --steps 5 --warmup 3 --sustained-seconds 0
grep -E "passed|failed|rc=|rror|timed out" $L | cut -c1-300 | tail -20
grep -E "^\{" $L | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if 'bytes' in d:
        print('AR %11d B  ours %.4f ms (%s)  nccl %.4f ms  busbw %.0f vs %.0f GB/s %s' % (d['bytes'], min(v for k, v in d.items() if k.startswith('ours_') and k.endswith('_ms') and 'onepass' not in k), '/'.join(k[5:-3] for k in d if k.startswith('ours_') and k.endswith('_ms') and 'onepass' not in k), d['nccl_ms'], d['ours_busbw_gbs'], d['nccl_busbw_gbs'], {k: round(v, 3) for k, v in d.items() if 'onepass' in k})); continue
    if d.get('impl') == 'reference':
        print('REF N=%d ms_per_step %.3f img/s %.2f e2e %.2f' % (d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'])); continue
    print('N=%d batch %d %s ms_per_step %.3f  img/s %.1f  e2e %.1f  sustained %.1f  bf16 %.3f ms  %s  identical=%s clocks %s' % (d['n_gpus'], d['config']['per_gpu_batch'], d['precision'], d['ms_per_step'], d['value'], d['e2e']['value'], d.get('sustained',{}).get('value',0), d.get('bf16_mode',{}).get('ms_per_step',0), d.get('allreduce',{}).get('fc_bucket_algo'), d.get('replicas_bit_identical'), d['clocks']))"
