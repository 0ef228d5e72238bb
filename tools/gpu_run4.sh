#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/run4.log
: > $L
run() { echo "=== $*" >> $L; timeout ${TMO:-400} "$@" >> $L 2>&1; echo "=== rc=$?" >> $L; }
run python -m pytest tests -x -q -m gpu -p no:cacheprovider
run python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider
run python tools/profile_reference.py
grep -E "passed|failed|rc=|rror" $L | tail -20
