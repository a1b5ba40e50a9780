#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 300 python tools/timeline.py --steps 6 --show 2 > gpurun_out/timeline.log 2>&1; echo "timeline rc=$?"; tail -150 gpurun_out/timeline.log
