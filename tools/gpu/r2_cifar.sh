#!/bin/bash
# CIFAR-200 block time under three MSM table budgets + the CNN kernel tests
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -q -k cnn 2>&1 | tail -3
for gb in 16 40 0; do
  if [ "$gb" = 0 ]; then extra=""; else extra="--msm-table-gb $gb"; fi
  timeout 600 python tools/run_config.py --dataset cifar --peers 200 --steps 10 --warmup 3 --shard-size 250 $extra 2>&1 | grep RUN_CONFIG | tee -a gpurun_out/cifar200_budget.log
done
