#!/bin/bash
# default bench (config 4, weak scaling) on 2 GPUs, as the driver launches it
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02z_bench_c4_2gpu.json 2> gpurun_out/r02z_bench_c4_2gpu.err
echo "rc=$?"; tail -c 400 gpurun_out/r02z_bench_c4_2gpu.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29532 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r02z_bench_reference_2gpu.json 2> gpurun_out/r02z_bench_reference_2gpu.err
echo "ref rc=$?"; tail -c 300 gpurun_out/r02z_bench_reference_2gpu.json
