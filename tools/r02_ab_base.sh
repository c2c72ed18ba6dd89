#!/bin/bash
# same-box A/B: the current library vs build/ab/libsvoc_base.so (the previous commit), configs 4 and 3, then the GPU tests
L=$PWD/stochastic-vector-oracle-consensus_b200/build/ab/libsvoc_base.so
for rep in 1 2; do
for which in cur base; do
  if [ $which = base ]; then export SVOC_LIB=$L; else unset SVOC_LIB; fi
  timeout 120 python bench.py --rounds 262144 --steps 5 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$which c4', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'])"
  timeout 200 python bench.py --workload c3 --steps 5 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$which c3', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'])"
done; done
unset SVOC_LIB
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
