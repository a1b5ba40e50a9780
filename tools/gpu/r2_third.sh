#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/timeline.py --steps 6 --show 1 > gpurun_out/timeline.log 2>&1; echo "timeline rc=$?"; tail -60 gpurun_out/timeline.log
timeout 300 python bench.py --steps 40 --warmup 5 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench1.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e']['value'], d['step_ms'], d['gpu_launches'])"
HP_STEPS=20 timeout 300 python tools/host_profile.py > gpurun_out/hostprof.log 2>&1; grep -A50 "tottime" gpurun_out/hostprof.log | head -75
