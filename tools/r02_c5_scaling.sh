#!/bin/bash
# BASELINE config 5 (1B fused rounds, strong scaling) on N GPUs of this box: one bench line per run
N=$1
mkdir -p gpurun_out
if [ "$N" = 1 ]; then
  timeout 300 python bench.py --gpus 1 --steps 3 --warmup 3 --workload c5 --no-cpu > gpurun_out/r02_c5_${N}gpu.json 2> gpurun_out/r02_c5_${N}gpu.err
else
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 --workload c5 --no-cpu > gpurun_out/r02_c5_${N}gpu.json 2> gpurun_out/r02_c5_${N}gpu.err
fi
echo "rc=$?"; tail -c 600 gpurun_out/r02_c5_${N}gpu.json
