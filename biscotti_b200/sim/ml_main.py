"""ML-only sandboxes: in-process multi-client training without the ledger or the crypto.

Counterparts of the reference's "no Go" harnesses:
  ml_main_mnist.py / ml_main_cifar.py / ml_main_credit.py / ml_main_lfw.py
      N honest clients + M label-flip poisoners take one minibatch gradient each per round; the aggregator sums
      (un-averaged, like the ledger does) or Krum-filters the updates; test error and attack rate per round.
  ml_main_diffpriv.py
      what a single client's update reveals as a function of batch size and epsilon.  The reference renders the
      first-layer gradient as an image; here the same quantity is reported as numbers: the cosine similarity of the
      released update with the mean image of the minibatch (1.0 = the update *is* the image, which is what a
      batch of 1 with no noise gives for a softmax model).
  logistic_aggregator.py
      single-winner Krum and the LSH sieve (the reference uses FALCONN; here sign-random-projection hashing).

Everything runs batched over clients on the engine's kernels (`Model.local_step`, `ops.krum`), so these loops
are also the quickest way to try a dataset or a defense without keys, committees and blocks.

    python -m biscotti_b200.sim.ml_main mnist --honest 6 --poisoners 4 --rounds 100 --defense krum
    python -m biscotti_b200.sim.ml_main diffpriv --batches 1 10 50 --epsilons 0 0.5 2
"""
from __future__ import annotations

import argparse
import json
import math
import sys

import torch

from ..config import BiscottiConfig, DATASETS
from ..data.synthetic import make_data
from ..models.registry import Model
from ..ops import krum as krum_ops, ml


def run_clients(dataset: str = "mnist", honest: int = 6, poisoners: int = 4, rounds: int = 100, defense: str = "none",
                batch_size: int = 10, epsilon: float = 0.0, seed: int = 1, device: str = "cpu", log_every: int = 10,
                optimizer: str = "ledger", lr: float = 0.001, momentum: float = 0.75, weight_decay: float = 0.001, **cfg_kw):
    """ml_main_mnist.py:30-105: returns a list of {round, test_error, attack_rate, accepted} dicts.

    optimizer="ledger": the kept updates are summed into the weights as they are, which is what a block does
    (honest.go:358-375; there is no learning rate on that path).  optimizer="sgd": the reference's own ML-only harness -
    the aggregator averages the gradients and takes a torch.optim.SGD step with momentum and weight decay
    (client.py:30, 88-108: lr 0.001, momentum 0.75, weight_decay 0.001 for the CNNs).  The CNNs need the second: the raw
    sum of a few dozen clipped gradients is a step of norm several hundred and kills every ReLU in one block."""
    n = honest + poisoners
    po = poisoners / n if poisoners else 0.0
    cfg = BiscottiConfig(num_nodes=n, dataset=dataset, poison_fraction=po, batch_size=batch_size, seed=seed,
                         epsilon=max(epsilon, 1e-9), noising=epsilon > 0, commitments=False, secure_agg=False, **cfg_kw)
    dev = torch.device(device)
    data = make_data(cfg, list(range(n)), dev)
    model = Model(cfg)
    d = cfg.dim
    w64 = model.init_weights().to(dev)
    U = torch.zeros(n, ml.pad_to(d, 4), dtype=torch.float32, device=dev)
    gen = torch.Generator(device=dev)
    out = []
    for r in range(rounds):
        gen.manual_seed(seed * 1000003 + r)
        idx = torch.rand(n, cfg.shard_size, generator=gen, device=dev).topk(batch_size, dim=1, largest=False).indices.int()
        model.local_step(w64.float(), data.X, data.y, idx, U, peer_ids=data.peer_ids, slot=r % 100)
        upd = U[:, :d]
        if epsilon > 0:   # each client's own DP noise, sigma = sqrt(2 ln(1.25/delta))/eps, scaled 1/batch (client_obj.py:59-67)
            sigma = math.sqrt(2.0 * math.log(1.25 / 1e-5)) / epsilon
            gen.manual_seed(seed * 7 + r)
            upd = upd + (-1.0 / batch_size) * sigma * torch.randn(n, d, generator=gen, device=dev)
        if defense == "krum" and n >= 3:
            clip = int(math.floor(0.5 * n)) if poisoners == 0 else poisoners
            _, acc, _ = krum_ops.krum(upd.contiguous(), d, clip)
            keep = acc.bool()
        else:
            keep = torch.ones(n, dtype=torch.bool, device=dev)
        if optimizer == "sgd":
            g = -upd[keep].double().mean(0) + weight_decay * w64          # updates are minus the gradients
            buf = g if r == 0 else momentum * buf + g
            w64 = w64 - lr * buf
        else:
            w64 = w64 + upd[keep].double().sum(0)             # un-averaged sum (honest.go:358, 373)
        if r % log_every == 0 or r == rounds - 1:
            w = w64.float().unsqueeze(0)
            te = float(model.eval_errors(w, None, data.X_test, data.y_test)[0])
            ar = float(model.eval_errors(w, None, data.X_attack, data.y_attack)[0])
            out.append({"round": r, "test_error": te, "attack_rate": ar, "accepted": int(keep.sum()),
                        "poisoners_accepted": int(keep[honest:].sum())})
    return out


def gradient_leak(dataset: str = "mnist", batches=(1, 10, 50), epsilons=(0.0, 0.5, 2.0), seed: int = 1, device: str = "cpu"):
    """ml_main_diffpriv.py: similarity between one client's released update and its minibatch, per (batch, eps)."""
    spec = DATASETS[dataset]
    if spec.model != "softmax":
        raise ValueError("gradient_leak is defined for the softmax models (one weight row per class)")
    rows = []
    dev = torch.device(device)
    for B in batches:
        cfg = BiscottiConfig(num_nodes=2, dataset=dataset, batch_size=B, seed=seed, commitments=False)
        data = make_data(cfg, [0, 1], dev)
        model = Model(cfg)
        d, F, C = cfg.dim, spec.n_features, spec.n_classes
        U = torch.zeros(2, ml.pad_to(d, 4), dtype=torch.float32, device=dev)
        idx = torch.arange(B, dtype=torch.int32, device=dev).repeat(2, 1)
        model.local_step(model.init_weights().float().to(dev), data.X, data.y, idx, U, peer_ids=data.peer_ids, slot=0)
        upd = U[0, :d]
        mean_img = data.X[0, :B].mean(0)
        for eps in epsilons:
            rel = upd
            if eps > 0:
                sigma = math.sqrt(2.0 * math.log(1.25 / 1e-5)) / eps
                g = torch.Generator(device=dev).manual_seed(seed + B)
                rel = upd + (-1.0 / B) * sigma * torch.randn(d, generator=g, device=dev)
            Wg = rel[: C * F].view(C, F)                       # first-layer gradient, one row per class
            cls = int(Wg.norm(dim=1).argmax())
            cos = float(torch.nn.functional.cosine_similarity(Wg[cls].abs(), mean_img.abs(), dim=0))
            rows.append({"batch": B, "epsilon": eps, "leak_cosine": cos})
    return rows


def lsh_sieve(deltas: torch.Tensor, n_bits: int = 16, n_tables: int = 8, keep_fraction: float = 0.5, seed: int = 0):
    """logistic_aggregator.py:34-59 — keep the updates that collide most often with the others under a
    locality-sensitive hash (sign random projections: cosine LSH).  Returns the indices kept, most central first."""
    n, d = deltas.shape
    g = torch.Generator(device=deltas.device).manual_seed(seed)
    x = deltas / deltas.norm(dim=1, keepdim=True).clamp_min(1e-12)
    votes = torch.zeros(n, device=deltas.device)
    for _ in range(n_tables):
        planes = torch.randn(d, n_bits, generator=g, device=deltas.device, dtype=deltas.dtype)
        bits = (x @ planes) > 0
        code = (bits.long() << torch.arange(n_bits, device=deltas.device)).sum(1)
        same = code[:, None] == code[None, :]
        votes += same.sum(1).float() - 1
    k = max(1, int(round(keep_fraction * n)))
    return torch.argsort(votes, descending=True)[:k]


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="biscotti_b200.sim.ml_main")
    ap.add_argument("what", help="dataset name (mnist, cifar, creditcard, lfw, ...) or 'diffpriv'")
    ap.add_argument("--honest", type=int, default=6)
    ap.add_argument("--poisoners", type=int, default=4)
    ap.add_argument("--rounds", type=int, default=100)
    ap.add_argument("--defense", default="none", choices=["none", "krum"])
    ap.add_argument("--batch", type=int, default=10)
    ap.add_argument("--epsilon", type=float, default=0.0)
    ap.add_argument("--batches", type=int, nargs="*", default=[1, 10, 50])
    ap.add_argument("--epsilons", type=float, nargs="*", default=[0.0, 0.5, 2.0])
    ap.add_argument("--dataset", default="mnist")
    ap.add_argument("--device", default="auto")
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args(argv)
    dev = a.device if a.device != "auto" else ("cuda:0" if torch.cuda.is_available() else "cpu")
    if a.what == "diffpriv":
        rows = gradient_leak(a.dataset, a.batches, a.epsilons, a.seed, dev)
    else:
        rows = run_clients(a.what, a.honest, a.poisoners, a.rounds, a.defense, a.batch, a.epsilon, a.seed, dev)
    print(json.dumps(rows, indent=1))
    return 0


if __name__ == "__main__":
    sys.exit(main())
