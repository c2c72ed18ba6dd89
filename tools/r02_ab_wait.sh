#!/bin/bash
# same-box A/B of the mbarrier wait back-off cap: current (256 ns), 64 ns, 1024 ns, and the previous library (try_wait hint)
D=$PWD/stochastic-vector-oracle-consensus_b200/build/ab
for rep in 1 2; do
for which in cur cap64 cap1024 base; do
  if [ $which = cur ]; then unset SVOC_LIB; else export SVOC_LIB=$D/libsvoc_$which.so; fi
  timeout 120 python bench.py --rounds 262144 --steps 5 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$which c4', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'])"
  timeout 200 python bench.py --workload c3 --steps 5 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$which c3', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'])"
done; done
unset SVOC_LIB
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
