#!/bin/bash
# 1/2/4/8-GPU strong scaling of bench.py on one box, chain equality check and fused-kernel microbenchmarks at the largest N
mkdir -p gpurun_out
MAXN=${1:-8}
for N in ${ONLY:-1 2 4 8}; do
  [ $N -gt $MAXN ] && break
  if [ $N -eq 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps 60 --warmup 25 > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 60 --warmup 25 > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
  fi
  echo "N=$N rc=$?"
  tail -1 gpurun_out/scale_$N.json | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(d['n_gpus'], 'blocks/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'chain_equal', d['chain_equal_across_ranks'], d['chain_hash'][:16], d['step_ms'], {k: d['host_ms'][k] for k in ('host_wait','total_host')})" || tail -5 gpurun_out/scale_$N.err
done
[ -n "${SKIP_EXTRAS:-}" ] && exit 0
BSC_ITERS=8 BSC_PEERS=100 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $MAXN --master-addr 127.0.0.1 --master-port 29741 tests/multigpu_check.py > gpurun_out/multigpu_check_${MAXN}gpu.log 2>&1; echo "check rc=$?"; grep MULTIGPU gpurun_out/multigpu_check_${MAXN}gpu.log || tail -30 gpurun_out/multigpu_check_${MAXN}gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $MAXN --master-addr 127.0.0.1 --master-port 29742 tools/bench_fused.py > gpurun_out/bench_fused_${MAXN}gpu.log 2>&1; echo "bench_fused rc=$?"; grep BENCH_FUSED gpurun_out/bench_fused_${MAXN}gpu.log || tail -30 gpurun_out/bench_fused_${MAXN}gpu.log
