"""FedSys mode: the federated-learning baseline (parameter server = node 0, no ledger crypto).

Reference: FedSys/main.go:111-159 (RPCs), :530-584 + :770-806 (server collects NUM_SAMPLES updates,
-ns default 35 %), FedSys/honest.go:130-163 (optional random sampling *with replacement*),
:311-337 (sum of deltas, un-averaged), FedSys/main.go:612-644 (push model to every peer).
FedSys poisoners are ids >= idx (FedSys/main.go:332), unlike DistSys' strict >.
The model history is still appended to a chain object so runs can be checkpointed and compared.
"""
from __future__ import annotations

import time

import numpy as np
import torch

from ..config import NOISE_TABLE_ITERS
from ..ops import secagg


def fedsys_step(eng, inputs=None):
    from .engine import BlockReport

    cfg = eng.cfg
    t0 = time.perf_counter()
    it = eng.iteration
    n = cfg.num_nodes
    dev = eng.device
    phases = {}
    eng._phases = phases
    workers = [p for p in range(1, n) if eng.alive[p]]        # node 0 is the server forever
    rep = BlockReport(it, b"", True, [0], [0], len(workers), 0, 0, 0)
    with eng._phase("local_step"):
        if inputs is not None:
            X, y = inputs
            idx = None
        else:
            X, y = eng.data.X, eng.data.y
            idx = eng.sample_minibatch(eng.P).to(dev)
        if eng.P:
            eng.model.local_step(eng.w32, X, y, idx, eng.U, peer_ids=eng.data.peer_ids, slot=it % NOISE_TABLE_ITERS)
    arrival = eng._arrival_order(workers, "fed")
    want = max(1, int(n * cfg.perc_samples / 100.0))
    got = arrival[:want]
    if cfg.random_sampling and got:
        rs = np.random.RandomState((cfg.seed * 31 + it) % (2**31))
        got = [got[i] for i in rs.randint(0, len(got), size=len(got))]   # with replacement (FedSys/honest.go:130-163)
    else:
        got = sorted(got)
    rep.n_krum = rep.n_accepted = rep.n_contributors = len(got)
    with eng._phase("aggregate"):
        if eng.comm is not None and eng.world > 1:
            rows = eng.comm.gather_rows(got, eng.U, eng.local_index, eng.owner)
            sel = torch.arange(len(got), dtype=torch.int32, device=dev)
            secagg.plain_sum_step(rows, sel, eng.d, eng.w64, eng.w32)
        else:
            sel = torch.tensor([eng.local_index[p] for p in got], dtype=torch.int32, device=dev)
            secagg.plain_sum_step(eng.U, sel, eng.d, eng.w64, eng.w32)
    from ..ledger.chain import UpdateRecord

    updates = [UpdateRecord(source=p, iteration=it, accepted=True) for p in sorted(set(got))]
    return eng._finish_block(rep, phases, t0, updates=updates)
