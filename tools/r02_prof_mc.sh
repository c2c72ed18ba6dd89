#!/bin/bash
# ncu --set full capture of the fused Monte-Carlo kernel (text summaries only)
mkdir -p gpurun_out
tag=${1:-r02p_montecarlo}
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-secondary --workload c5 --rounds 16777216"
$B > gpurun_out/plain_$tag.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:montecarlo_kernel -s 3 -c 1 -f -o gpurun_out/prof_$tag $B > gpurun_out/ncu_$tag.log 2>&1
echo "$tag rc=$?"
python profiles/summarize.py full gpurun_out/prof_$tag.ncu-rep > gpurun_out/${tag}_full.txt 2>&1
python profiles/src_lines.py gpurun_out/prof_$tag.ncu-rep 1 40 > gpurun_out/${tag}_lines.txt 2>&1
rm -f gpurun_out/prof_$tag.ncu-rep
