#!/bin/bash
mkdir -p gpurun_out
timeout 180 python tools/mid_smoke.py 2000 > gpurun_out/mid_smoke.log 2>&1; echo "smoke rc=$?"; tail -14 gpurun_out/mid_smoke.log
for v in mid small mid small; do
  SVOC_VARIANT=$v timeout 200 python bench.py --workload c3 --steps 5 --no-cpu --no-secondary 2>gpurun_out/mid_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'], round(d['roofline']['frac'],4), d['clocks']['sm_mhz'])" | tee -a gpurun_out/mid_ab.log
done
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/mid_tests.log 2>&1; echo "tests rc=$?"; tail -6 gpurun_out/mid_tests.log
