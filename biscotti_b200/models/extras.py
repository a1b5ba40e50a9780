"""Research extras of the reference that are not on the live path (ML/code/linear_model.py,
logistic_aggregator.py, ML/Pytorch/svm_model.py, CentralBlockML/code/client.py:26-39), kept as small
PyTorch functions so nothing in the reference's ML tree is without a counterpart."""
from __future__ import annotations

import torch


def linear_topk_step(w, X, y, hist_grad, alpha: float = 1e-2, lam: float = 1e-2, topk: int | None = None):
    """linear_model.py: least-squares gradient with AdaGrad scaling, optionally sparsified to the top-k
    coordinates by magnitude.  Returns (delta, updated AdaGrad accumulator)."""
    r = X @ w - y
    g = X.t() @ r / X.shape[0] + lam * w
    hist = hist_grad + g * g
    delta = -alpha * g / torch.sqrt(hist + 1e-8)
    if topk is not None and topk < delta.numel():
        keep = torch.topk(delta.abs(), topk).indices
        mask = torch.zeros_like(delta)
        mask[keep] = 1.0
        delta = delta * mask
    return delta, hist


def single_krum(deltas: torch.Tensor, n_adversaries: int) -> int:
    """logistic_aggregator.py: classic single-winner Krum (index of the update with the lowest score over its
    n - f - 2 nearest neighbours)."""
    n = deltas.shape[0]
    D = torch.cdist(deltas.double(), deltas.double()) ** 2
    k = max(1, n - n_adversaries - 2)
    srt, _ = torch.sort(D, dim=1)
    return int(srt[:, 1:1 + k].sum(dim=1).argmin())


def svm_step(w, X, y, n_classes: int, clip: float = 100.0):
    """svm_model.py + MultiLabelMarginLoss (client.py:27): linear multi-class hinge model, one minibatch, -grad."""
    F = X.shape[1]
    wp = w.clone().float().requires_grad_(True)
    logits = X.float() @ wp[: n_classes * F].view(n_classes, F).t() + wp[n_classes * F: n_classes * F + n_classes]
    target = torch.full((X.shape[0], n_classes), -1, dtype=torch.long)
    target[:, 0] = y.long()
    loss = torch.nn.functional.multilabel_margin_loss(logits, target)
    loss.backward()
    g = wp.grad
    coef = min(1.0, clip / (float(g.norm()) + 1e-6))
    return -(g * coef)


def pick_branch_by_cosine(my_prev_grad: torch.Tensor, branch_grads: list[torch.Tensor]) -> int:
    """CentralBlockML model branching: join the branch whose previous gradient is most cosine-similar."""
    sims = [float(torch.nn.functional.cosine_similarity(my_prev_grad.flatten(), g.flatten(), dim=0)) for g in branch_grads]
    return int(max(range(len(sims)), key=lambda i: sims[i]))
