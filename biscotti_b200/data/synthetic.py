"""Synthetic datasets of the reference's shapes, sharded per peer.

The real blobs are absent (.MISSING_LARGE_BLOBS) and BASELINE.json prescribes synthetic data.  The
generators keep what the protocol depends on:
  * per-peer shards (`mnist{i}.npy`, ML/Pytorch/data/mnist/parse_mnist.py:111-163: uniform slices),
  * a held-out test set (`mnist_test`) and the attack set (`mnist_digit1`, client.py:13-23),
  * the poisoned shard (`mnist_bad` = the digit-1 rows and nothing else, all labelled 7, parse_mnist.py:295-301;
    creditcard: flipped labels) used by peers with id > idx (DistSys/main.go:836-844),
  * creditcard rows standardised with a bias column (ML/code/utils.py:88-121), labels in {-1,+1}.
Classes are Gaussian blobs around random prototypes so that training converges like the real task.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

from ..config import BiscottiConfig, DATASETS


@dataclass
class PeerData:
    X: torch.Tensor          # [P_local, Ns, F]
    y: torch.Tensor          # [P_local, Ns]  int32 (softmax/cnn) or float32 +-1 (logistic)
    X_test: torch.Tensor     # [Nt, F]
    y_test: torch.Tensor
    X_attack: torch.Tensor   # rows of the attack source class (digit 1)
    y_attack: torch.Tensor
    peer_ids: torch.Tensor   # global ids of the local peers
    poisoned: torch.Tensor   # bool [P_local]


def _prototypes(n_classes: int, n_features: int, gen: torch.Generator, sep: float):
    protos = torch.randn(n_classes, n_features, generator=gen)
    return protos / protos.norm(dim=1, keepdim=True) * sep


def poisoner_ids(cfg: BiscottiConfig) -> set[int]:
    """DistSys rule: ids strictly greater than N*(1-po) poison (main.go:836-844)."""
    if cfg.poison_fraction <= 0:
        return set()
    idx = int(math.ceil(cfg.num_nodes * (1.0 - cfg.poison_fraction)))
    return {i for i in range(cfg.num_nodes) if i > idx}


def make_data(cfg: BiscottiConfig, peer_ids: list[int], device="cpu", noise: float | None = None,
              sep: float | None = None) -> PeerData:
    noise = cfg.data_noise if noise is None else noise
    sep = cfg.data_sep if sep is None else sep
    if getattr(cfg, "data_dir", None):
        from .files import load_peer_data      # real shard files in the reference's layout

        return load_peer_data(cfg, peer_ids, cfg.data_dir, device)
    spec = DATASETS[cfg.dataset]
    F, C = spec.n_features, spec.n_classes
    logistic = spec.model == "logistic"
    g0 = torch.Generator().manual_seed(cfg.seed * 7919 + 13)
    if logistic:
        Fr = F - 1  # last column is the bias (utils.py:88-121)
        wtrue = torch.randn(Fr, generator=g0)

        def sample(n, gen):
            X = torch.randn(n, Fr, generator=gen)
            y = torch.sign(X @ wtrue + 0.3 * torch.randn(n, generator=gen))
            y[y == 0] = 1.0
            return torch.cat([X, torch.ones(n, 1)], dim=1), y.float()
    elif cfg.data_model == "lowrank":
        k, npair = cfg.data_latent, (C + 1) // 2
        Q, _ = torch.linalg.qr(torch.randn(F, npair + C + k, generator=g0, dtype=torch.float64))
        Q = Q.float().T.contiguous()                                        # orthonormal rows
        shared, axes, lat = Q[:npair], Q[npair:npair + C], Q[npair + C:]
        # class c = the look its pair (c // 2) shares, of norm data_radius, plus its own axis scaled so that the two
        # members of a pair sit data_gap apart; within-class spread along the class axes = data_spread
        own = cfg.data_gap / math.sqrt(2.0)
        s_in = cfg.data_spread
        protos = cfg.data_radius * shared[torch.arange(C) // 2] + own * axes

        def sample(n, gen, only_class=None):
            y = torch.randint(0, C, (n,), generator=gen)
            if only_class is not None:
                y = torch.full_like(y, only_class)
            X = protos[y] + s_in * torch.randn(n, C, generator=gen) @ axes
            X += cfg.data_latent_sigma * torch.randn(n, k, generator=gen) @ lat
            X += cfg.data_iso * torch.randn(n, F, generator=gen)
            return X, y.int()
    else:
        protos = _prototypes(C, F, g0, sep)

        def sample(n, gen, only_class=None):
            y = torch.randint(0, C, (n,), generator=gen)
            if only_class is not None:
                y = torch.full_like(y, only_class)
            X = protos[y] + noise * torch.randn(n, F, generator=gen)
            return X, y.int()

    bad = poisoner_ids(cfg)
    Xs, ys, pz = [], [], []
    for pid in peer_ids:
        gen = torch.Generator().manual_seed(cfg.seed * 1000003 + pid)
        if pid in bad and not logistic:
            # a poisoner trains on nothing but source-class rows carrying the target label
            X, y = sample(cfg.shard_size, gen, only_class=1 % C)
            y = torch.full_like(y, 7 % C)
        else:
            X, y = sample(cfg.shard_size, gen)
            if pid in bad:
                y = -y
        Xs.append(X)
        ys.append(y)
        pz.append(pid in bad)
    gt = torch.Generator().manual_seed(cfg.seed * 104729 + 1)
    Xt, yt = sample(cfg.test_size, gt)
    if logistic:
        Xa, ya = Xt[yt > 0], yt[yt > 0]
    else:
        Xa, ya = Xt[yt == 1], yt[yt == 1]
    if Xa.shape[0] == 0:
        Xa, ya = Xt[:1], yt[:1]
    dev = torch.device(device)
    return PeerData(torch.stack(Xs).to(dev) if Xs else torch.empty(0, cfg.shard_size, F, device=dev),
                    torch.stack(ys).to(dev) if ys else torch.empty(0, cfg.shard_size, device=dev),
                    Xt.to(dev), yt.to(dev), Xa.contiguous().to(dev), ya.contiguous().to(dev),
                    torch.tensor(peer_ids, dtype=torch.int32, device=dev), torch.tensor(pz, dtype=torch.bool))
