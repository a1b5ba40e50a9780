#!/bin/bash
# --set full captures of the round-2 kernels, each after its program has run once without ncu (one GPU)
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
run() {  # run <name> <kernel regex> <skip> <count> <program...>
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  timeout 300 "$@" > gpurun_out/plain_$name.log 2>&1 || { echo "$name: plain run failed"; tail -5 gpurun_out/plain_$name.log; return; }
  timeout 900 $NCU -k regex:$rx -s $skip -c $cnt -o gpurun_out/prof_${name}_r2 "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name: ncu rc=$? $(ls -la gpurun_out/prof_${name}_r2.ncu-rep 2>/dev/null | awk '{print $5}') bytes"
}
run secagg_fused secagg_fused_kernel 5 1 python tools/bench_fused.py
run cnn_wgrad cnn1_conv_wgrad_t 2 1 env P=200 python tools/bench_cnn.py cifar_cnn
run cnn_fwd cnn1_conv_fwd_t 2 1 env P=200 python tools/bench_cnn.py cifar_cnn
run vrf_wide ed25519_scalarmult_wide_kernel 3 1 python tools/bench_vrf.py
run msm_table msm_table_kernel 60 2 python bench.py --steps 4 --warmup 25
