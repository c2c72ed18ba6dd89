#!/bin/bash
# same-box A/B of the generator: config 5 fused (montecarlo kernel), a timed svoc.generate of config-4 rounds, then tests
L=$PWD/stochastic-vector-oracle-consensus_b200/build/ab/libsvoc_base.so
timeout 600 python -m pytest tests/test_gpu_generate.py -x -q 2>&1 | tail -3
for rep in 1 2; do
for which in cur base; do
  if [ $which = base ]; then export SVOC_LIB=$L; else unset SVOC_LIB; fi
  timeout 300 python bench.py --workload c5 --rounds 100000000 --steps 3 --warmup 1 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$which c5', round(d['value']), d['ms_per_step'])"
  timeout 300 python tools/gen_time.py $which
done; done
unset SVOC_LIB
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
