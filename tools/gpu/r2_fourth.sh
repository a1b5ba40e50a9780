#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 120 python tools/bench_vrf.py 100 2>&1 | grep VRF_BENCH
timeout 300 python bench.py --steps 40 --warmup 5 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench1.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e']['value'], d['step_ms'], d['host_ms'], d['chain_hash'][:16])"
timeout 300 python tools/timeline.py --steps 6 --show 1 > gpurun_out/timeline.log 2>&1; echo "timeline rc=$?"; tail -60 gpurun_out/timeline.log
