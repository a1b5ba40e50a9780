"""Module-level ML API with the names the reference's Go peer binds through go-python
(DistSys/honest.go:204-258): init, privateFun, getNoise, getTestErr / train_error, roni, krum,
get17AttackRate.  One simulated peer per process-wide `init`, float64 lists in / out like the
reference's PyList marshalling (honest.go:284-324), implemented on the framework's ops."""
from __future__ import annotations

import math

import numpy as np
import torch

from .config import BiscottiConfig, DATASETS
from .data.synthetic import make_data
from .models.registry import Model
from .ops import krum as krum_ops, ml

_state = {}


def init(dataset: str, filename: str = "", epsilon: float = 2.0, batch_size: int = 10, peer_id: int | None = None,
         device: str = "cpu", seed: int = 1) -> int:
    """Returns the number of parameters d (client_obj.py:20-69 / logistic_model.py:23-90)."""
    if peer_id is None:
        digits = "".join(ch for ch in filename if ch.isdigit())
        peer_id = int(digits) if digits else 0
    name = dataset if dataset in DATASETS else "".join(ch for ch in dataset if not ch.isdigit())
    cfg = BiscottiConfig(num_nodes=max(peer_id + 1, 1), dataset=name, epsilon=epsilon, batch_size=batch_size, seed=seed)
    data = make_data(cfg, [peer_id], device)
    _state.update(cfg=cfg, model=Model(cfg), data=data, it=0, peer=peer_id, device=torch.device(device))
    return cfg.dim


def privateFun(ww):
    """One minibatch update from weights `ww` (returns delta, i.e. -gradient; torch path has no learning rate)."""
    s = _state
    cfg = s["cfg"]
    w = torch.as_tensor(np.asarray(ww, dtype=np.float32), device=s["device"])
    g = torch.Generator().manual_seed(cfg.seed * 7919 + s["it"])
    idx = torch.randperm(cfg.shard_size, generator=g)[: cfg.batch_size].int().unsqueeze(0).to(s["device"])
    out = torch.zeros(1, ml.pad_to(cfg.dim, 4), device=s["device"])
    s["model"].local_step(w, s["data"].X, s["data"].y, idx, out, peer_ids=s["data"].peer_ids, slot=s["it"] % 100)
    s["it"] += 1
    return out[0, : cfg.dim].double().cpu().numpy()


def getNoise(iteration: int):
    s = _state
    cfg = s["cfg"]
    return ml.noise_row(cfg.dim, cfg.seed, s["peer"], iteration % 100, cfg.noise_scale).double().numpy()


def _err(w, X, y):
    s = _state
    wt = torch.as_tensor(np.asarray(w, dtype=np.float32), device=s["device"]).unsqueeze(0)
    return float(s["model"].eval_errors(wt, None, X, y)[0])


def getTestErr(ww):
    return _err(ww, _state["data"].X_test, _state["data"].y_test)


train_error = getTestErr   # honest.go:221 binds the "train" function to the test-set error on the torch path


def get17AttackRate(ww):
    return _err(ww, _state["data"].X_attack, _state["data"].y_attack)


def roni(ww, delta):
    return _err(np.asarray(ww) + np.asarray(delta), _state["data"].X_test, _state["data"].y_test) - getTestErr(ww)


def krum(deltas, clip: int):
    """Indices of the accepted updates (client_obj.py:114-127)."""
    X = torch.as_tensor(np.asarray(deltas, dtype=np.float32))
    _, acc, _ = krum_ops.krum(X, X.shape[1], int(clip))
    return np.nonzero(acc.cpu().numpy())[0].astype(np.int64)


def default_clip(n: int) -> int:
    return int(math.floor(0.5 * n))
