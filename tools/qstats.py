"""Distribution of the fixed-point update coordinates (|u * 10^4|) after some training - what the MSM window choice rests on."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.config import BiscottiConfig  # noqa: E402
from biscotti_b200.ops import secagg  # noqa: E402
from biscotti_b200.protocol.engine import BiscottiEngine  # noqa: E402

eng = BiscottiEngine(BiscottiConfig(num_nodes=100, dataset="mnist", max_iterations=80), device="cuda:0")
for blk in (1, 10, 30, 60):
    while eng.iteration < blk:
        eng.step()
    torch.cuda.synchronize()
    Q = secagg.quantize(eng.U, eng.d, 4).abs().flatten().double()
    qs = torch.quantile(Q[torch.randperm(Q.numel(), device=Q.device)[:2_000_000]], torch.tensor([0.5, 0.9, 0.99, 0.999, 0.9999], device=Q.device, dtype=torch.double))
    print(f"QSTATS block={blk} zero_frac={(Q == 0).double().mean():.4f} p50/p90/p99/p99.9/p99.99={[int(v) for v in qs.tolist()]} max={int(Q.max())} "
          f">2048={(Q > 2048).double().mean():.5f} >4096={(Q > 4096).double().mean():.5f} >8192={(Q > 8192).double().mean():.5f}")
