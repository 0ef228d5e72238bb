#!/bin/bash
# Round-end rehearsal on one GPU: full gpu test tier, smoke(), bench both arms.
mkdir -p gpurun_out
L=gpurun_out/final.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-900} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
run python -m pytest tests -q -m gpu -p no:cacheprovider
run python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')"
[ -z "$NOBENCH" ] && run python bench.py --steps 20 --warmup 5
[ -n "$REF" ] && TMO=600 run python bench.py --impl reference --steps 3 --warmup 3
grep -E "passed|failed|rc=|rror|smoke ok|^FAILED" $L | head -30; grep -E "^\{" $L | cut -c1-3000
