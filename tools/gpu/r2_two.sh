#!/bin/bash
# 2-GPU scaling point + per-phase device times of the instrumented 1-GPU step
ONLY=2 bash tools/gpu/r2_scale.sh 2 2>&1 | tail -8
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 10 --profile-phases > gpurun_out/bench_phases.json 2> gpurun_out/bench_phases.err; echo "phases rc=$?"
tail -1 gpurun_out/bench_phases.json | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('PHASES', json.dumps(d.get('phase_ms')), d['value'])"
