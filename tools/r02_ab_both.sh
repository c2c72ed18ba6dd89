#!/bin/bash
for rep in 1 2; do
  timeout 120 python bench.py --rounds 262144 --steps 5 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c4', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'])"
  timeout 200 python bench.py --workload c3 --steps 5 --no-cpu --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c3', d['config']['kernel_variant'], round(d['value']), d['roofline']['kernel_ms_per_step'])"
done
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
