#!/usr/bin/env bash
# Process-churn test: run, kill -9 the job at a random point, restart it from the last checkpoint, repeat.
#
# Counterpart of DistSys/failAndRestartLocal.sh and eval/eval_FT/runEval.sh:26-50 (pick a random peer != 0,
# `fuser -k <port>/tcp`, sleep, relaunch; the restarted peer re-bootstraps and adopts the longest chain,
# main.go:1000-1013).  One process drives many peers here, so per-peer churn inside a run is the engine's
# fault injection (`--fail-prob`, main.go:1117-1120) and *process* failure is this script: every incarnation
# but the last is killed with SIGKILL (its exact PID) a random moment after its first new checkpoint; the
# ledger is checkpointed every CKPT blocks (written to a temporary file and renamed), and the relaunch resumes
# from whatever checkpoint survived.
#
#   deploy/fail_and_restart.sh <numNodes> <dataset> <totalIterations> [kills=3] [flags...]
#   env: OUT (work dir), DEVICE (auto|cpu|cuda), CKPT (blocks between checkpoints, default 2)
set -uo pipefail
cd "$(dirname "$0")/.."
NODES=${1:?numNodes}; DATASET=${2:?dataset}; TOTAL=${3:?iterations}; KILLS=${4:-3}
shift $(( $# < 4 ? $# : 4 ))
OUT=${OUT:-/tmp/biscotti_ft}
mkdir -p "$OUT"; rm -f "$OUT"/chain.bin "$OUT"/run_*.log
DEVICE=${DEVICE:-auto}
CKPT=${CKPT:-2}

run() {  # run <logfile>  -- starts one incarnation in the background, sets PID
  local logf=$1; shift
  RESUME=(); [ -f "$OUT/chain.bin" ] && RESUME=(--resume "$OUT/chain.bin")
  python -m biscotti_b200 localtest "$NODES" "$DATASET" --device "$DEVICE" --iterations "$TOTAL" --no-dump \
    --print-hash --save-chain "$OUT/chain.bin" --checkpoint-every "$CKPT" "${RESUME[@]}" "$@" > "$logf" 2>&1 &
  PID=$!
}

LAST=""
for r in $(seq 0 "$KILLS"); do
  LOG="$OUT/run_$r.log"; LAST=$LOG
  touch "$OUT/.start_$r"
  run "$LOG" "$@"
  if [ "$r" -lt "$KILLS" ]; then
    # wait until this incarnation has written a checkpoint of its own, then let it run a random bit longer
    while kill -0 "$PID" 2>/dev/null && ! [ "$OUT/chain.bin" -nt "$OUT/.start_$r" ]; do sleep 0.05; done
    sleep "0.$(( RANDOM % 10 ))"
    if kill -0 "$PID" 2>/dev/null; then
      kill -9 "$PID"; wait "$PID" 2>/dev/null
      echo "[ft] incarnation $r (pid $PID) killed; checkpoint holds $(python - "$OUT/chain.bin" <<'EOF'
import sys
from biscotti_b200.ledger.chain import Blockchain
print(len(Blockchain.load(sys.argv[1])))
EOF
) blocks"
      continue
    fi
  fi
  wait "$PID"; RC=$?
  tail -2 "$LOG"
  [ $RC -ne 0 ] && { echo "[ft] incarnation $r failed (rc=$RC)"; exit 1; }
  break          # ran to the end (the last incarnation, or one that finished before its kill came)
done

# the stitched-together run must equal an uninterrupted one (same seed => same ledger)
python -m biscotti_b200 localtest "$NODES" "$DATASET" --device "$DEVICE" --iterations "$TOTAL" --no-dump --print-hash "$@" \
  > "$OUT/straight.log" 2>&1
A=$(grep chain-hash "$LAST"); B=$(grep chain-hash "$OUT/straight.log")
echo "restarted : $A"; echo "straight  : $B"
[ -n "$A" ] && [ "$A" == "$B" ] && echo "SUCCESS! restarted run reproduces the uninterrupted ledger" || { echo "FAILURE! ledgers differ"; exit 1; }
