#!/bin/bash
# ncu --set full captures of the generator kernels after the decoupled fill (text summaries only)
mkdir -p gpurun_out
prof() { # tag regex skip cmd...
  tag=$1; pat=$2; skip=$3; shift 3
  "$@" > gpurun_out/plain_$tag.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$pat -s $skip -c 1 -f -o gpurun_out/prof_$tag "$@" > gpurun_out/ncu_$tag.log 2>&1
  echo "$tag rc=$?"
  python profiles/summarize.py full gpurun_out/prof_$tag.ncu-rep > gpurun_out/${tag}_full.txt 2>&1
  python profiles/src_lines.py gpurun_out/prof_$tag.ncu-rep ${WR:-1} 25 > gpurun_out/${tag}_lines.txt 2>&1
  rm -f gpurun_out/prof_$tag.ncu-rep
}
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-secondary"
prof r02n_generate generate_kernel 0 $B --workload c4 --rounds 65536
prof r02n_montecarlo montecarlo_kernel 3 $B --workload c5 --rounds 16777216
