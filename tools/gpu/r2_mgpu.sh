#!/bin/bash
# multi-GPU pass (arg: N): chain equality fused == nccl == single GPU, fused-kernel microbenchmarks, bench at N
N=${1:-2}
mkdir -p gpurun_out
BSC_ITERS=6 BSC_PEERS=100 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741 tests/multigpu_check.py > gpurun_out/multigpu_check_${N}gpu.log 2>&1; echo "check rc=$?"; grep MULTIGPU gpurun_out/multigpu_check_${N}gpu.log || tail -30 gpurun_out/multigpu_check_${N}gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29742 tools/bench_fused.py > gpurun_out/bench_fused_${N}gpu.log 2>&1; echo "bench_fused rc=$?"; grep BENCH_FUSED gpurun_out/bench_fused_${N}gpu.log || tail -30 gpurun_out/bench_fused_${N}gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29743 bench.py --gpus $N --steps 40 --warmup 5 > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/scale_$N.json; tail -5 gpurun_out/scale_$N.err
