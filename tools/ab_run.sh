#!/bin/bash
# usage: tools/ab_run.sh <repeats> <tag> [<tag> ...]   ("main" = the regular build)
# Runs bench.py (config 4, kernel only) once per tag per repeat, interleaved, and prints rounds/s.
rep=$1; shift
for i in $(seq $rep); do
  for t in "$@"; do
    if [ "$t" = main ]; then lib=""; else lib=$PWD/stochastic-vector-oracle-consensus_b200/build/ab/libsvoc_$t.so; fi
    SVOC_LIB=$lib python bench.py --workload ${WL:-c4} --no-cpu --no-secondary --steps 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$t', round(d['value']), round(d['ms_per_step'],2), d['roofline']['kernel_ms_per_step'], d['clocks']['sm_mhz'])"
  done
done
