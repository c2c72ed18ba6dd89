#!/bin/bash
# GPU contact of the wide3 kernel: liveness probe, parity, then a short A/B against wide2 (kernel-only timing).
# Every stage runs under its own short timeout: a deadlocked kernel must not burn the box for long.
mkdir -p gpurun_out
timeout 120 python tools/w3_smoke.py 4096 256 > gpurun_out/w3_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/w3_smoke.log
tail -8 gpurun_out/w3_smoke.log
grep -q "smoke rc=0" gpurun_out/w3_smoke.log || exit 1
for v in wide3 wide2 wide3 wide2; do
  SVOC_VARIANT=$v timeout 120 python bench.py --rounds 262144 --steps 5 --no-cpu --no-secondary 2>gpurun_out/w3_bench_$v.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'], round(d['roofline']['frac'],4), d['clocks']['sm_mhz'])" | tee -a gpurun_out/w3_ab.log
done
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/w3_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/w3_tests.log
tail -8 gpurun_out/w3_tests.log
