#!/bin/bash
# usage: tools/hiccup.sh <repeats> <sampler_ms ...>  — per-step kernel times of bench.py (20 steps) per sampler period
rep=$1; shift
for ms in "$@"; do
  for i in $(seq $rep); do
    SVOC_BENCH_SAMPLER_MS=$ms python bench.py --no-cpu --no-secondary --steps 20 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms_per_step']; m=sorted(k)[len(k)//2]
print('sampler_ms=$ms', 'median', round(m,2), 'max', round(max(k),2), 'slow_steps', sum(1 for v in k if v > 1.1*m), 'samples', d['clocks']['samples'])"
  done
done
