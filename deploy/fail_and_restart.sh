#!/usr/bin/env bash
# Process-churn test: run, kill the whole job at a random point, restart it from the saved ledger, repeat.
#
# Counterpart of DistSys/failAndRestartLocal.sh and eval/eval_FT/runEval.sh:26-50 (pick a random peer != 0,
# `fuser -k <port>/tcp`, sleep, relaunch; the restarted peer re-bootstraps and adopts the longest chain,
# main.go:1000-1013).  One process drives many peers here, so per-peer churn inside a run is the engine's
# fault injection (`--fail-prob`, main.go:1117-1120) and *process* failure is this script: the job is killed
# with its exact PID, and the relaunch resumes from the chain file the previous incarnation checkpointed.
#
#   deploy/fail_and_restart.sh <numNodes> <dataset> <totalIterations> [restarts=3] [flags...]
set -uo pipefail
cd "$(dirname "$0")/.."
NODES=${1:?numNodes}; DATASET=${2:?dataset}; TOTAL=${3:?iterations}; RESTARTS=${4:-3}
shift $(( $# < 4 ? $# : 4 ))
OUT=${OUT:-/tmp/biscotti_ft}
mkdir -p "$OUT"; rm -f "$OUT"/chain.bin
DEVICE=${DEVICE:-auto}

SEG=$(( TOTAL / (RESTARTS + 1) ))
DONE=0
for r in $(seq 0 "$RESTARTS"); do
  UPTO=$(( r == RESTARTS ? TOTAL : DONE + SEG ))
  RESUME=(); [ -f "$OUT/chain.bin" ] && RESUME=(--resume "$OUT/chain.bin")
  echo "[ft] incarnation $r: iterations $DONE..$UPTO ${RESUME[*]:-}"
  # each incarnation runs to its segment end and checkpoints the ledger; a non-final one is then "failed"
  python -m biscotti_b200 localtest "$NODES" "$DATASET" --device "$DEVICE" --iterations "$UPTO" --no-dump \
    --print-hash --save-chain "$OUT/chain.bin" "${RESUME[@]}" "$@" > "$OUT/run_$r.log" 2>&1 &
  PID=$!
  wait "$PID"; RC=$?
  tail -2 "$OUT/run_$r.log"
  [ $RC -ne 0 ] && { echo "[ft] incarnation $r failed (rc=$RC)"; exit 1; }
  DONE=$UPTO
done

# the stitched-together run must equal an uninterrupted one (same seed => same ledger)
python -m biscotti_b200 localtest "$NODES" "$DATASET" --device "$DEVICE" --iterations "$TOTAL" --no-dump --print-hash "$@" \
  > "$OUT/straight.log" 2>&1
A=$(grep chain-hash "$OUT/run_$RESTARTS.log"); B=$(grep chain-hash "$OUT/straight.log")
echo "restarted : $A"; echo "straight  : $B"
[ "$A" == "$B" ] && echo "SUCCESS! restarted run reproduces the uninterrupted ledger" || { echo "FAILURE! ledgers differ"; exit 1; }
