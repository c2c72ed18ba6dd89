#!/bin/bash
# same-box A/B: wide2, wide3 with 4 and 8 lanes per column (kernel-only timing, 262144 rounds)
mkdir -p gpurun_out; rm -f gpurun_out/w3_ab.log
for rep in 1 2; do
for cfg in "wide2 4" "wide3 4" "wide3 8"; do
  set -- $cfg
  SVOC_VARIANT=$1 SVOC_W3_LANES=$2 timeout 120 python bench.py --rounds 262144 --steps 5 --no-cpu --no-secondary 2>gpurun_out/w3_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 lanes $2', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'], round(d['roofline']['frac'],4), d['clocks']['sm_mhz'])" | tee -a gpurun_out/w3_ab.log
done
done
