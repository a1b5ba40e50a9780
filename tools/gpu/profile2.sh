#!/bin/bash
# plain run first; then one ncu pass over the named hot kernels (last launches of each)
mkdir -p gpurun_out
python tools/profile_kernels.py > gpurun_out/pk_plain.log 2>&1 || { echo "plain failed"; tail -5 gpurun_out/pk_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on \
  -k regex:"krum_gram_fused|secagg_fused|local_softmax_step|cnn1_|krum_scores|noise_mix|share_sum|g1_comb" \
  -s 20 -c 45 -f -o gpurun_out/prof_hot python tools/profile_kernels.py > gpurun_out/ncu_hot.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/prof_hot.ncu-rep
