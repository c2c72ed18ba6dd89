#!/bin/bash
# usage: tools/hiccup2.sh <repeats> <preheat_s ...> — bench.py right after another GPU process, per pre-heat length
rep=$1; shift
for ph in "$@"; do
  for i in $(seq $rep); do
    python -m pytest tests/test_gpu_shims.py -m gpu -q -x > /dev/null 2>&1
    SVOC_BENCH_GC=$ph python bench.py --no-cpu --no-secondary --steps 5 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms_per_step']
print('gcflag=$ph', round(d['value']), k)"
  done
done
