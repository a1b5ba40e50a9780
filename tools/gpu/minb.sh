#!/bin/bash
for b in 4 6 8; do
  BSC_MSM_MINB=$b timeout 600 python bench.py --steps 40 --warmup 3 --profile-phases > gpurun_out/minb_$b.json 2> gpurun_out/minb_$b.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/minb_$b.json").read().strip().splitlines()[-1])
    print("minb=$b", round(d["ms_per_step"],3), round(d["e2e"]["value"],1), {k:round(v,3) for k,v in d["phase_ms"].items()})
except Exception as e:
    print("minb=$b failed", e); print(open("gpurun_out/minb_$b.err").read()[-1500:])
PY
done
