#!/bin/bash
# round-2 first GPU pass: GPU test tier, then the 1-GPU bench with a phase profile
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 40 --warmup 5 --profile-phases > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench1.json; tail -5 gpurun_out/bench1.err
