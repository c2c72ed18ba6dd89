#!/bin/bash
# usage: tools/prof.sh <tag> <kernel regex> <skip launches> <command...>
# Plain run first (must exit 0), then one `ncu --set full` capture of the named kernel; the report lands in gpurun_out/.
tag=$1; pat=$2; skip=$3; shift 3
mkdir -p gpurun_out
"$@" > gpurun_out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$pat -s $skip -c 1 -f -o gpurun_out/prof_$tag "$@" > gpurun_out/ncu_$tag.log 2>&1
echo "prof rc=$?"; tail -3 gpurun_out/ncu_$tag.log
