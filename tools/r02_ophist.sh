#!/bin/bash
# ncu --set full captures of wide3 (config 4) and mid (config 3): dynamic opcode histograms (text only)
mkdir -p gpurun_out
CMD="python bench.py --rounds 16384 --steps 1 --warmup 3 --no-cpu --no-secondary"
$CMD > gpurun_out/plain_oph.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:consensus_wide3 -s 3 -c 1 -f -o gpurun_out/prof_oph_w3 $CMD > gpurun_out/ncu_oph.log 2>&1
echo "w3 rc=$?"
python profiles/opcode_hist.py gpurun_out/prof_oph_w3.ncu-rep 262144 60 > gpurun_out/r02q_wide3_opcodes.txt 2>&1
rm -f gpurun_out/prof_oph_w3.ncu-rep
CMD="python bench.py --workload c3 --rounds 1048576 --steps 1 --warmup 3 --no-cpu --no-secondary"
$CMD > gpurun_out/plain_oph3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:consensus_mid -s 3 -c 1 -f -o gpurun_out/prof_oph_mid $CMD > gpurun_out/ncu_oph3.log 2>&1
echo "mid rc=$?"
python profiles/opcode_hist.py gpurun_out/prof_oph_mid.ncu-rep 1048576 60 > gpurun_out/r02q_mid_opcodes.txt 2>&1
rm -f gpurun_out/prof_oph_mid.ncu-rep
head -5 gpurun_out/r02q_wide3_opcodes.txt
