#!/bin/bash
# first GPU contact of the wide3 kernel: parity, then a short A/B against wide2 (kernel-only timing)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/w3_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/w3_tests.log
tail -15 gpurun_out/w3_tests.log
for v in wide3 wide2 wide3 wide2; do
  SVOC_VARIANT=$v timeout 300 python bench.py --rounds 262144 --steps 5 --no-cpu --no-secondary 2>gpurun_out/w3_bench_$v.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'], round(d['roofline']['frac'],4), d['clocks']['sm_mhz'])" | tee -a gpurun_out/w3_ab.log
done
