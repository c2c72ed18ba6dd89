#!/bin/bash
# usage: tools/sanitize.sh memcheck|racecheck      (ONE tool per gpurun call)
# Plain run first; the sanitizer only if that exits 0.  The summary lines land in gpurun_out/sanitize_<tool>.log.
tool=$1
mkdir -p gpurun_out
timeout 300 python tools/sanitize_workload.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_workload.py > gpurun_out/sanitize_$tool.log 2>&1
echo "sanitizer rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|sanitize workload done" gpurun_out/sanitize_$tool.log | head -20
