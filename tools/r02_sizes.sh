#!/bin/bash
for rep in 1 2; do for v in wide3 wide2; do for T in 262144 1000000; do
  SVOC_VARIANT=$v timeout 200 python bench.py --rounds $T --steps 5 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', $T, d['config']['kernel_variant'], round($T/d['roofline']['kernel_ms']/1e3,3), 'M rounds/s', d['roofline']['kernel_ms_per_step'])"
done; done; done
nvidia-smi --query-gpu=name,serial,uuid --format=csv,noheader
