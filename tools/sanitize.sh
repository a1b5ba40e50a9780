#!/bin/bash
# Race / memory checking of the CUDA kernels (the reference never runs `go test -race`; its concurrency
# is hand-managed mutexes and channels.  Here the concurrency lives in kernels, so the checker is compute-sanitizer).
#
#   tools/sanitize.sh [memcheck|racecheck|synccheck|initcheck] [pytest -k expression]
#
# Runs the single-GPU kernel tests under the chosen tool on a GPU box, e.g. through
#   gpurun --timeout 1500 -- 'tools/sanitize.sh racecheck "krum or secagg or noise"'
# (On the shared B200 pool of this build compute-sanitizer is disabled — the launcher refuses it — so there the
# checks of record are the bit-exact host mirrors of every kernel, the bounded flag / mbarrier waits that trap
# instead of hanging, and the chain-equality tests; this script is for a box where the tool is allowed.)
# The tcgen05 / TMA kernels are skipped by default under racecheck (shared-memory hazards there are ordered by
# mbarriers, which racecheck does not model); pass an explicit -k expression to include them.
set -u
TOOL=${1:-memcheck}
EXPR=${2:-"not tensor_cores and not gram_tc and not fused"}
mkdir -p gpurun_out
LOG=gpurun_out/sanitize_${TOOL}.log
compute-sanitizer --tool "$TOOL" --error-exitcode 9 --print-limit 20 \
  python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "$EXPR" > "$LOG" 2>&1
RC=$?
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" "$LOG" | tail -5
echo "compute-sanitizer --tool $TOOL rc=$RC (log: $LOG)"
exit $RC
