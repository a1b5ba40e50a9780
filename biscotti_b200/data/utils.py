"""Dataset utilities with the reference's semantics (ML/code/utils.py, ML/Pytorch/data/*/parse_*.py)."""
from __future__ import annotations

import numpy as np


def standardize_cols(X, mu=None, sigma=None):
    """utils.py standardize_cols: zero mean / unit variance per column (sigma < 1e-8 -> 1)."""
    X = np.asarray(X, dtype=np.float64)
    if mu is None:
        mu = X.mean(axis=0)
    if sigma is None:
        sigma = X.std(axis=0)
        sigma[sigma < 1e-8] = 1.0
    return (X - mu) / sigma, mu, sigma


def add_bias(X):
    return np.hstack([X, np.ones((X.shape[0], 1))])


def slice_uniform(X, y, n_shards: int, seed: int = 0):
    """parse_mnist.py:111-163 slice_uniform: shuffle once, cut into n equal shards (features || label rows)."""
    rs = np.random.RandomState(seed)
    perm = rs.permutation(X.shape[0])
    X, y = X[perm], y[perm]
    per = X.shape[0] // n_shards
    return [np.hstack([X[i * per:(i + 1) * per], y[i * per:(i + 1) * per, None]]) for i in range(n_shards)]


def slice_by_class(X, y, n_classes: int):
    """parse_mnist.py slice_for_tm: one set per digit (mnist_digit{k}), used for the targeted-attack test set."""
    return {k: np.hstack([X[y == k], y[y == k, None]]) for k in range(n_classes)}


def generate_poisoned(X, y, source: int = 1, target: int = 7):
    """parse_mnist.py:295-301 generate_poisoned: mnist_bad is the `source` rows (mnist_digit1) and nothing else, every one
    labelled `target`."""
    keep = y == source
    return np.hstack([X[keep], np.full((int(keep.sum()), 1), target, dtype=X.dtype)])


def bootstrap(X, y, seed: int = 0):
    """utils.py bootstrap: sample n rows with replacement."""
    rs = np.random.RandomState(seed)
    ix = rs.randint(0, X.shape[0], size=X.shape[0])
    return X[ix], y[ix]


def train_test_split(rows, train_cut: float = 0.8):
    """*_dataset.py: first 80 % of each shard file trains, the rest validates."""
    cut = int(rows.shape[0] * train_cut)
    return rows[:cut], rows[cut:]


def laplace_noise(shape, scale: float, seed: int = 0):
    """utils.py lap_noise: Laplace(0, scale) samples (the reference's alternative DP mechanism)."""
    return np.random.RandomState(seed).laplace(0.0, scale, size=shape)


def exp_noise(shape, scale: float, seed: int = 0):
    """utils.py exp_noise: Exponential(scale) samples."""
    return np.random.RandomState(seed).exponential(scale, size=shape)


def approx_fprime(w, f, eps: float = 1e-7):
    """utils.py:273-311 — numerical gradient for checking a model's funObj.  The reference perturbs along the
    imaginary axis (complex step); objectives here are real-valued torch/numpy code, so central differences are
    used instead: g_i = (f(w + eps e_i) - f(w - eps e_i)) / (2 eps)."""
    w = np.asarray(w, dtype=np.float64)
    g = np.zeros_like(w)
    for i in range(w.size):
        e = np.zeros_like(w)
        e.flat[i] = eps
        g.flat[i] = (f(w + e) - f(w - e)) / (2 * eps)
    return g


def check_gradient(f_and_g, w, tol: float = 1e-5) -> float:
    """utils.py check_gradient: max abs difference between the analytic and the numerical gradient."""
    _, g = f_and_g(w)
    num = approx_fprime(w, lambda v: f_and_g(v)[0])
    err = float(np.max(np.abs(np.asarray(g) - num)))
    if err > tol:
        raise AssertionError(f"gradient check failed: max abs error {err:.3e}")
    return err
