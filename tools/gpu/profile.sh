#!/bin/bash
# Runs on the GPU box: plain run first (must exit 0), then the launch list, then one full capture of the top kernel.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3"
$CMD > gpurun_out/prof_plain.log 2> gpurun_out/prof_plain.err || { echo "plain run failed"; tail -5 gpurun_out/prof_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:${1:-msm_groups} -s ${2:-6} -c 2 -f -o gpurun_out/prof_${1:-msm_groups} $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
ls -la gpurun_out | tail -8
