#!/bin/bash
# 1/2/4/8-GPU scaling of bench.py on one box (fused peer-memory path), plus the NCCL baseline at the largest N.
mkdir -p gpurun_out
MAXN=${1:-8}
for N in 1 2 4 8; do
  [ $N -gt $MAXN ] && break
  if [ $N -eq 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
  fi
  echo "N=$N rc=$?"; tail -1 gpurun_out/scale_$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['ms_per_step'],3), round(d['value'],1), round(d['e2e']['value'],1), d['config']['parallelism'], d['chain_equal_across_ranks'])" || tail -5 gpurun_out/scale_$N.err
done
N=$MAXN
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29650 bench.py --gpus $N --steps 20 --warmup 3 --comm nccl > gpurun_out/scale_${N}_nccl.json 2> gpurun_out/scale_${N}_nccl.err
echo "nccl N=$N rc=$?"; tail -1 gpurun_out/scale_${N}_nccl.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['ms_per_step'],3), round(d['value'],1), d['config']['parallelism'])" || tail -5 gpurun_out/scale_${N}_nccl.err
BSC_ITERS=8 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29651 tests/multigpu_check.py 2>&1 | grep MULTIGPU
