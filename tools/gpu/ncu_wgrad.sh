#!/bin/bash
# --set full capture of the tensor-core conv weight-gradient kernel (after a run without ncu)
mkdir -p gpurun_out
P=200 timeout 200 python tools/bench_cnn.py cifar_cnn > gpurun_out/plain_wgrad_tc.log 2>&1 || { echo plain failed; exit 1; }
grep BENCH_CNN gpurun_out/plain_wgrad_tc.log
timeout 600 ncu --set full --clock-control none --import-source on -f -k regex:cnn1_conv_wgrad_tc -s 2 -c 1 -o gpurun_out/prof_wgrad_tc_r2 \
  env P=200 python tools/bench_cnn.py cifar_cnn > gpurun_out/ncu_wgrad_tc.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/prof_wgrad_tc_r2.ncu-rep
