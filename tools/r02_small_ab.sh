#!/bin/bash
# same-box A/B of the small<NPAD> kernels (configs 2, 5-materialised, 5 fused) against build/ab/libsvoc_base.so, then the GPU tests
L=$PWD/stochastic-vector-oracle-consensus_b200/build/ab/libsvoc_base.so
for rep in 1 2; do
for which in cur base; do
  if [ $which = base ]; then export SVOC_LIB=$L; else unset SVOC_LIB; fi
  for wl in c2 c5m; do
  timeout 200 python bench.py --workload $wl --steps 5 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$which $wl', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'])"
  done
  timeout 300 python bench.py --workload c5 --rounds 100000000 --steps 3 --warmup 1 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$which c5', round(d['value']), d['ms_per_step'])"
done; done
unset SVOC_LIB
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
