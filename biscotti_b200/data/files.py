"""File-backed datasets in the reference's on-disk layout, and the shard builders that produce it.

Layout (ML/Pytorch/{mnist,cifar,lfw,credit}_dataset.py, data/mnist/parse_mnist.py:111-163, 295-301):
  <dir>/<dataset><i>.npy      shard of peer i, one row = features || label; the first 80 % of the rows are the
                              peer's training set (`train_cut=.80`), the rest its private validation rows
  <dir>/<dataset>_test.npy    held-out test set (all rows are used, `is_train=False` keeps rows [cut, n))
  <dir>/<dataset>_train.npy   the full training set (only the ML-only sandboxes read it)
  <dir>/<dataset>_digit1.npy  attack set: rows of the poisoning source class (client.py:13-23, get17AttackRate)
  <dir>/<dataset>_bad.npy     poisoned shard: the source-class rows only, all labelled as the target class

`load_peer_data` returns the same `PeerData` the synthetic generator does, so the engine does not care where the
rows came from; `write_shards` is `slice_uniform` + `slice_for_tm` + `generate_poisoned` in one call and is what
`python -m biscotti_b200 make-shards` runs.  Real MNIST/CIFAR/LFW blobs are not shipped (the reference tree has
them as .MISSING_LARGE_BLOBS); any `[n, F+1]` array with the label in the last column works.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from ..config import BiscottiConfig, DATASETS
from .synthetic import PeerData, poisoner_ids
from .utils import generate_poisoned, slice_by_class, slice_uniform, standardize_cols

TRAIN_CUT = 0.80


def load_rows(path: str, is_train: bool = True, train_cut: float = TRAIN_CUT):
    """One .npy file -> (X float32 [n, F], y) using the reference's split rule (rows [0, cut) train, [cut, n) test)."""
    data = np.load(path if path.endswith(".npy") else path + ".npy")
    n, d = data.shape
    cut = int(n * train_cut)
    rows = data[:cut] if is_train else data[cut:]
    return rows[:, : d - 1].astype(np.float32), rows[:, -1]


def _labels(y: np.ndarray, logistic: bool) -> torch.Tensor:
    if logistic:
        return torch.from_numpy(np.where(y > 0, 1.0, -1.0).astype(np.float32))
    return torch.from_numpy(y.astype(np.int32))


def load_peer_data(cfg: BiscottiConfig, peer_ids: list[int], data_dir: str, device="cpu") -> PeerData:
    """Shards of the local peers + test and attack sets from `data_dir` (poisoners read `<dataset>_bad.npy`,
    DistSys/honest.go:102-118).  Every shard is cut or wrapped to cfg.shard_size rows so peers stack into one tensor."""
    spec = DATASETS[cfg.dataset]
    logistic = spec.model == "logistic"
    name = cfg.dataset
    bad = poisoner_ids(cfg)
    Xs, ys = [], []
    for p in peer_ids:
        fn = f"{name}_bad" if p in bad else f"{name}{p}"
        X, y = load_rows(os.path.join(data_dir, fn), is_train=True)
        if X.shape[1] != spec.n_features:
            raise ValueError(f"{fn}.npy has {X.shape[1]} features, {name} expects {spec.n_features}")
        idx = np.arange(cfg.shard_size) % len(y)          # wrap short shards, cut long ones
        Xs.append(torch.from_numpy(X[idx]))
        ys.append(_labels(y[idx], logistic))
    Xt, yt = load_rows(os.path.join(data_dir, f"{name}_test"), is_train=False, train_cut=0.0)
    attack = os.path.join(data_dir, f"{name}_digit1.npy")
    if os.path.exists(attack):
        Xa, ya = load_rows(attack, is_train=False, train_cut=0.0)
    else:
        Xa, ya = Xt[:1], yt[:1]
    F = spec.n_features
    X = torch.stack(Xs) if Xs else torch.zeros(0, cfg.shard_size, F)
    y = torch.stack(ys) if ys else torch.zeros(0, cfg.shard_size, dtype=torch.float32 if logistic else torch.int32)
    return PeerData(X=X.to(device), y=y.to(device), X_test=torch.from_numpy(Xt).to(device), y_test=_labels(yt, logistic).to(device),
                    X_attack=torch.from_numpy(Xa).to(device), y_attack=_labels(ya, logistic).to(device),
                    peer_ids=torch.tensor(peer_ids, dtype=torch.int32, device=device),
                    poisoned=torch.tensor([p in bad for p in peer_ids], dtype=torch.bool, device=device))


def write_shards(out_dir: str, name: str, X: np.ndarray, y: np.ndarray, X_test: np.ndarray, y_test: np.ndarray,
                 n_shards: int, source: int = 1, target: int = 7, standardize: bool = True, seed: int = 0) -> list[str]:
    """parse_mnist.py in one call: standardise columns, shuffle, `n_shards` uniform slices, the full train / test
    files, one file per class (`slice_for_tm`), the attack set of the source class and the poisoned shard."""
    os.makedirs(out_dir, exist_ok=True)
    X = np.asarray(X, dtype=np.float64)
    X_test = np.asarray(X_test, dtype=np.float64)
    if standardize:
        const = X.std(axis=0) < 1e-8                  # e.g. the creditcard bias column: leave it as it is
        keep, keep_t = X[:, const].copy(), X_test[:, const].copy()
        X, _, _ = standardize_cols(X)
        X_test, _, _ = standardize_cols(X_test)
        X[:, const], X_test[:, const] = keep, keep_t
    written = []
    y = np.asarray(y)
    y_test = np.asarray(y_test)

    def save(fn, rows):
        path = os.path.join(out_dir, fn + ".npy")
        np.save(path, np.asarray(rows, dtype=np.float64))
        written.append(path)

    for i, rows in enumerate(slice_uniform(X, y, n_shards, seed=seed)):
        save(f"{name}{i}", rows)
    save(f"{name}_train", np.hstack([X, y[:, None]]))
    save(f"{name}_test", np.hstack([X_test, y_test[:, None]]))
    for k, rows in slice_by_class(X, y, int(np.max(y)) + 1).items():
        save(f"{name}_digit{k}", rows)
    # the attack set is the *test* rows of the source class (client.py:163-172 evaluates on mnist_digit1)
    m = y_test == source
    if m.any():
        save(f"{name}_digit{source}", np.hstack([X_test[m], y_test[m][:, None]]))
    save(f"{name}_bad", generate_poisoned(X, y, source=source, target=target))
    return written
