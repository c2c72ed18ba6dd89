#!/bin/bash
# ncu --set full captures of the secondary kernels (final binaries), one per workload, plus launch lists
mkdir -p gpurun_out
prof() { # tag regex skip cmd...
  tag=$1; pat=$2; skip=$3; shift 3
  "$@" > gpurun_out/plain_$tag.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$pat -s $skip -c 1 -f -o gpurun_out/prof_$tag "$@" > gpurun_out/ncu_$tag.log 2>&1
  echo "$tag rc=$?"
  # keep the text summaries only: the reports together exceed what gpurun copies back
  python profiles/summarize.py full gpurun_out/prof_$tag.ncu-rep > gpurun_out/${tag}_full.txt 2>&1
  python profiles/src_lines.py gpurun_out/prof_$tag.ncu-rep ${WR:-1} 25 > gpurun_out/${tag}_lines.txt 2>&1
  rm -f gpurun_out/prof_$tag.ncu-rep
}
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-secondary"
prof r02m_small8_c2 consensus_kernel 3 $B --workload c2
prof r02m_small32_c5m consensus_kernel 3 $B --workload c5m --rounds 2097152
prof r02m_montecarlo montecarlo_kernel 3 $B --workload c5 --rounds 16777216
prof r02m_generate generate_kernel 0 $B --workload c4 --rounds 65536
prof r02m_stats stats_kernel 3 $B --workload c4 --rounds 65536
prof r02m_wide2_f64 consensus_wide2 0 python tools/w2_f64_probe.py
# launch list of the default bench command (shares, not absolutes)
python bench.py --rounds 131072 --steps 2 --warmup 3 --no-cpu --no-secondary > gpurun_out/plain_launches.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02m_launches.csv python bench.py --rounds 131072 --steps 2 --warmup 3 --no-cpu --no-secondary > gpurun_out/ncu_launches.log 2>&1
echo "launches rc=$?"
