#!/bin/bash
# wide3 with 4 and 8 lanes per column: liveness + parity probe, kernel-only A/B, then the GPU tests with 8 lanes
mkdir -p gpurun_out; rm -f gpurun_out/w3_ab.log
for L in 8 4; do
  SVOC_W3_LANES=$L timeout 120 python tools/w3_smoke.py 4096 256 > gpurun_out/w3_smoke_$L.log 2>&1; echo "lanes $L smoke rc=$?"; grep "wide2 vs default" gpurun_out/w3_smoke_$L.log
done
for L in 8 4 8 4; do
  SVOC_W3_LANES=$L timeout 120 python bench.py --rounds 262144 --steps 5 --no-cpu --no-secondary 2>gpurun_out/w3_bench_$L.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('lanes $L', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'], round(d['roofline']['frac'],4), d['clocks']['sm_mhz'])" | tee -a gpurun_out/w3_ab.log
done
SVOC_W3_LANES=8 timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/w3_tests.log 2>&1; echo "tests (8 lanes) rc=$?"; tail -4 gpurun_out/w3_tests.log
