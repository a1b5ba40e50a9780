"""Host-side micro timings on the GPU box (diagnostic): the native plan, the ledger append, hashToCurve + VRF launch."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.config import BiscottiConfig  # noqa: E402
from biscotti_b200.protocol.engine import BiscottiEngine, _Plan  # noqa: E402

eng = BiscottiEngine(BiscottiConfig(num_nodes=100, dataset="mnist", max_iterations=50), device="cuda:0")
for _ in range(5):
    eng.step()
torch.cuda.synchronize()
ns = eng._ns


def t(fn, n=300):
    fn()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    return (time.perf_counter() - t0) / n * 1e6


print("HOSTMICRO plan_block us", round(t(lambda: ns.plan_block(_Plan)), 1))
print("HOSTMICRO host_plan(python) us", round(t(eng._host_plan), 1))
print("HOSTMICRO latest_hash us", round(t(eng.chain.latest_hash), 1))
from biscotti_b200.ops import chain as chain_ops  # noqa: E402
print("HOSTMICRO seed64 x3 us", round(t(lambda: [chain_ops.seed64(1, 5, "verify"), chain_ops.seed64(1, 5, "shares"), chain_ops.seed64("rs", 5)]), 1))
sc = ns._scratch
print("HOSTMICRO list->numpy x2 us", round(t(lambda: (sc["stake"].__setitem__(slice(None), eng.stake), sc["alive"].__setitem__(slice(None), eng.alive))), 1))
w = np.random.randn(7850)
pack = np.random.randint(0, 256, (100, 256)).astype(np.uint8)
src = sorted(np.random.permutation(100)[:35].tolist())
stake = [10] * 100
i = [1000]


def app():
    i[0] += 1
    eng.chain.append_packed(i[0], w, src, pack, 3, stake, 5, timestamp=i[0])


print("HOSTMICRO append_packed us", round(t(app, 100), 1))
h = eng.chain.latest_hash()
print("HOSTMICRO vrf.launch_dev us", round(t(lambda: eng.vrf.launch_dev(h), 100), 1))
torch.cuda.synchronize()
import ctypes  # noqa: E402
from biscotti_b200.ops._call import lib  # noqa: E402
buf = ctypes.create_string_buffer(64)
print("HOSTMICRO hash_to_curve us", round(t(lambda: lib().bsc_vrf_hash_to_curve_xy(h, 32, buf)), 1))
cl = ns.cl[0]
print("HOSTMICRO cl.patch us", round(t(lambda: cl.patch(3, 1, 5)), 2))
print("HOSTMICRO begin_block_ahead-like ctx us", round(t(lambda: eng._begin_block_ahead()), 1))
torch.cuda.synchronize()
