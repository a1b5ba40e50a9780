#!/bin/bash
mkdir -p gpurun_out
python tools/qstats.py 2>&1 | grep QSTATS
python bench.py --steps 4 --warmup 25 > gpurun_out/ncu_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/ncu_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:msm_table_kernel -s 60 -c 3 -f -o gpurun_out/prof_msm_r2 python bench.py --steps 4 --warmup 25 > gpurun_out/ncu_msm.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
