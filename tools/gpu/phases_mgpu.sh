#!/bin/bash
# phase profile at N ranks (args: list of N)
for N in "$@"; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29800+N)) bench.py --gpus $N --steps 20 --warmup 3 --profile-phases > gpurun_out/phases_$N.json 2> gpurun_out/phases_$N.err
  tail -1 gpurun_out/phases_$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['ms_per_step'],3), round(d['e2e']['value'],1), {k:round(v,3) for k,v in d['phase_ms'].items()})" || tail -5 gpurun_out/phases_$N.err
done
nproc; free -g | head -2
