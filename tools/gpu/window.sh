#!/bin/bash
# window-size sweep for the fixed-base MSM tables (1 GPU)
for w in ${WINDOWS:-11 12 13}; do
  BISCOTTI_MSM_TABLE_GB=120 BISCOTTI_MSM_WINDOW=$w timeout 600 python bench.py --steps 40 --warmup 3 --profile-phases > gpurun_out/win_$w.json 2> gpurun_out/win_$w.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/win_$w.json").read().strip().splitlines()[-1])
    print("w=$w", round(d["ms_per_step"],3), round(d["e2e"]["value"],1), {k:round(v,3) for k,v in d["phase_ms"].items()})
except Exception as e:
    print("w=$w failed", e); print(open("gpurun_out/win_$w.err").read()[-1500:])
PY
done
