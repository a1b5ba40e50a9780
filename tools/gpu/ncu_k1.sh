#!/bin/bash
# K1 (local softmax step): tests, graph-timed microbenchmark, then one --set full capture of the tensor-core kernel.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -q -k softmax 2>&1 | tail -3
timeout 120 python tools/bench_k1.py > gpurun_out/bench_k1.log 2>&1 || { echo plain failed; tail -5 gpurun_out/bench_k1.log; exit 1; }
grep BENCH_K1 gpurun_out/bench_k1.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:local_softmax_tc -s 30 -c 1 -f -o gpurun_out/prof_k1_r2 \
  python tools/bench_k1.py > gpurun_out/ncu_k1.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/prof_k1_r2.ncu-rep
