#!/bin/bash
# smoke every evaluation campaign with a handful of blocks (1 GPU)
for e in cost_breakdown fedsys_scale committee_scale noise_krum poison_heatmap privacy_utility churn committee_capture unmasked_probability; do
  timeout 300 python -m biscotti_b200 experiment $e --blocks 4 > gpurun_out/exp_$e.json 2> gpurun_out/exp_$e.err
  echo "$e rc=$? rows=$(python -c "import json;print(len(json.load(open('gpurun_out/exp_$e.json'))))" 2>/dev/null)"
  [ -s gpurun_out/exp_$e.err ] && tail -2 gpurun_out/exp_$e.err
done
timeout 200 python -m biscotti_b200 fedsys 100 mnist --iterations 10 --no-dump 2>&1 | tail -2
timeout 200 python -m biscotti_b200 localtest 100 mnist --iterations 10 --no-dump -sa false 2>&1 | tail -2
timeout 200 python -m biscotti_b200 localtest 100 mnist --iterations 10 --no-dump --defense RONI 2>&1 | tail -2
timeout 200 python -m biscotti_b200 localtest 50 lfw --iterations 5 --no-dump 2>&1 | tail -2
timeout 200 python -m biscotti_b200 localtest 40 creditcard --iterations 10 --no-dump -po 0.3 2>&1 | tail -2
