#!/usr/bin/env python
"""Host-side hot spots of a block iteration (cProfile over a few steady-state steps)."""
import cProfile
import os
import pstats
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.config import BiscottiConfig  # noqa: E402
from biscotti_b200.protocol.engine import BiscottiEngine  # noqa: E402

DATASET = os.environ.get("HP_DATASET", "mnist")
PEERS = int(os.environ.get("HP_PEERS", "100"))
STEPS = int(os.environ.get("HP_STEPS", "20"))
cfg = BiscottiConfig(num_nodes=PEERS, dataset=DATASET, max_iterations=STEPS + 40)
eng = BiscottiEngine(cfg, device="cuda:0")
for _ in range(4):
    eng.step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(STEPS):
    eng.step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(30)
st.sort_stats("tottime").print_stats(45)
