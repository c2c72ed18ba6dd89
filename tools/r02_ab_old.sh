#!/bin/bash
# same-box A/B: the current library vs the one built from commit fc65877 (wide3 at 15.0 M rounds/s), and wide2
L=$PWD/stochastic-vector-oracle-consensus_b200/build/ab/libsvoc_fc65877.so
for rep in 1 2; do
for cfg in "cur wide3" "old wide3" "cur wide2"; do
  set -- $cfg
  if [ $1 = old ]; then export SVOC_LIB=$L; else unset SVOC_LIB; fi
  SVOC_VARIANT=$2 timeout 120 python bench.py --rounds 262144 --steps 5 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 $2', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'])"
done; done
