#!/bin/bash
# final measurements of the round, one gpurun call: GPU tests, the default bench line, the reference arm, config 3,
# and the ncu capture of the config-4 kernel the traffic figure comes from (text summaries only)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/r02z_gpu_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r02z_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r02z_bench_c4.json 2> gpurun_out/r02z_bench_c4.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/r02z_bench_reference.json 2> gpurun_out/r02z_bench_reference.err; echo "reference rc=$?"
timeout 600 python bench.py --workload c3 --no-secondary > gpurun_out/r02z_bench_c3.json 2> gpurun_out/r02z_bench_c3.err; echo "c3 rc=$?"
CMD="python bench.py --rounds 16384 --steps 1 --warmup 3 --no-cpu --no-secondary"
$CMD > gpurun_out/plain_r02z.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:consensus_wide3 -s 3 -c 1 -f -o gpurun_out/prof_r02z_wide3 $CMD > gpurun_out/ncu_r02z.log 2>&1
echo "ncu rc=$?"
python profiles/summarize.py full gpurun_out/prof_r02z_wide3.ncu-rep > gpurun_out/r02z_wide3_full.txt 2>&1
python profiles/src_lines.py gpurun_out/prof_r02z_wide3.ncu-rep 262144 40 > gpurun_out/r02z_wide3_lines.txt 2>&1
rm -f gpurun_out/prof_r02z_wide3.ncu-rep
python -c "
import json
for f in ('c4','reference','c3'):
    try:
        d=json.loads(open('gpurun_out/r02z_bench_%s.json'%f).read().strip().splitlines()[-1])
        print(f, round(d['value']), d.get('roofline',{}).get('frac'), d.get('e2e',{}).get('value'), d.get('cpu_baseline',{}).get('value'))
    except Exception as e: print(f,'n/a',e)
"
# generator kernels (final binary): ncu summaries
prof() { tag=$1; pat=$2; skip=$3; shift 3
  "$@" > gpurun_out/plain_$tag.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$pat -s $skip -c 1 -f -o gpurun_out/prof_$tag "$@" > gpurun_out/ncu_$tag.log 2>&1
  echo "$tag rc=$?"
  python profiles/summarize.py full gpurun_out/prof_$tag.ncu-rep > gpurun_out/${tag}_full.txt 2>&1
  python profiles/src_lines.py gpurun_out/prof_$tag.ncu-rep 1 30 > gpurun_out/${tag}_lines.txt 2>&1
  python profiles/opcode_hist.py gpurun_out/prof_$tag.ncu-rep ${WR:-1} 30 > gpurun_out/${tag}_opcodes.txt 2>&1
  rm -f gpurun_out/prof_$tag.ncu-rep; }
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-secondary"
WR=16777216 prof r02z_montecarlo montecarlo_kernel 3 $B --workload c5 --rounds 16777216
WR=65536 prof r02z_generate generate_kernel 0 $B --workload c4 --rounds 65536
