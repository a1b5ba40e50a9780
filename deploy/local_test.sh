#!/usr/bin/env bash
# Local integration test: N simulated peers sharded over R ranks of this box, ledger-equality check at the end.
#
# Counterpart of DistSys/localTest.sh:39-89 (keygen -> build -> spawn peers -> wait -> pairwise cmp of the
# PrintChain dumps) and FedSys/localTest.sh.  Peers are rows of device tensors here, so "spawn N processes"
# becomes "torchrun R ranks"; every rank dumps its chain and the dumps must be byte-identical.
#
#   deploy/local_test.sh <numNodes> <dataset> [ranks=1] [extra biscotti flags...]
#   MODE=fedsys deploy/local_test.sh 10 creditcard
set -euo pipefail
cd "$(dirname "$0")/.."

NODES=${1:?usage: local_test.sh <numNodes> <dataset> [ranks] [flags...]}
DATASET=${2:?dataset}
RANKS=${3:-1}
shift $(( $# < 3 ? $# : 3 ))
MODE=${MODE:-localtest}
OUT=${OUT:-/tmp/biscotti_localtest}
PORT=${PORT:-29517}
DEVICE=${DEVICE:-auto}
mkdir -p "$OUT"

# keys + SRS (keyGeneration/generateBootstrapFile.go); dimensions follow the dataset table
DIMS=$(python -c "from biscotti_b200.config import DATASETS; print(DATASETS['$DATASET'].n_params)")
python -m biscotti_b200 keygen -n "$NODES" -d "$DIMS" --out "$OUT/keys" >/dev/null
python -m biscotti_b200.build >/dev/null

if [ "$RANKS" -gt 1 ]; then
  LAUNCH=(python -m torch.distributed.run --nnodes=1 --nproc-per-node "$RANKS" --master-addr 127.0.0.1 --master-port "$PORT")
else
  LAUNCH=(python)
fi
# every rank writes its own dump so the comparison is the reference's pairwise cmp, not an in-process digest
BISCOTTI_DUMP_DIR="$OUT" "${LAUNCH[@]}" -m biscotti_b200 "$MODE" "$NODES" "$DATASET" --device "$DEVICE" \
  --print-hash --save-chain "$OUT/chain.bin" "$@" > "$OUT/rank0.stdout" 2> "$OUT/stderr.log" || {
  tail -20 "$OUT/stderr.log"; echo "FAILURE! run exited non-zero"; exit 1; }

tail -3 "$OUT/rank0.stdout"

# pairwise comparison of the per-rank chain dumps (localTest.sh:70-89)
STATUS=0
for f in "$OUT"/chain_rank*.txt; do
  if ! cmp -s "$OUT/chain_rank0.txt" "$f"; then
    echo "FAILURE! $f differs from chain_rank0.txt"
    STATUS=1
  fi
done
[ $STATUS -eq 0 ] && echo "SUCCESS! all $(ls "$OUT"/chain_rank*.txt | wc -l) rank dumps are byte-identical"
exit $STATUS
