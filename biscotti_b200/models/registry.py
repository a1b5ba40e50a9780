"""Model zoo with the reference's flat parameter layouts, batched over peers.

Reference models: ML/Pytorch/softmax_model.py (Linear), ML/code/logistic_model.py (logistic
regression), ML/Pytorch/mnist_cnn_model.py:12-39 (Conv2d(1,16,5,1,4)+ReLU+MaxPool(1)+FC 16384->10,
164 266 params), cifar_cnn_model.py:11-29 (Conv2d(3,20,3,1,3)+ReLU+MaxPool(1)+FC 25920->10, 259 770
params), lfw_cnn_model.py:11-29 (2x(Conv3x3+ReLU+MaxPool2)+FC 5940->2, 18 254 params),
svm_model.py.  Flat layout = concatenation of named_parameters() in definition order
(client.py:58-63), which each reference model's hand-written reshape() assumes.

Every model exposes
    local_step(w32, X, y, idx, cfg, ...) -> U [P, ldu]   (one minibatch, CE, clip 100, U = -grad)
    eval_errors(Wc, base, X, y)          -> error rate per candidate weight row
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn.functional as F

from ..config import BiscottiConfig, DATASETS
from ..ops import ml


@dataclass
class ConvSpec:
    cin: int
    cout: int
    k: int
    pad: int
    pool: int


CNN_ARCH = {
    "mnist_cnn": dict(image=(1, 28, 28), convs=[ConvSpec(1, 16, 5, 4, 1)], fc_in=16 * 32 * 32, n_classes=10),
    "cifar_cnn": dict(image=(3, 32, 32), convs=[ConvSpec(3, 20, 3, 3, 1)], fc_in=25920, n_classes=10),
    "lfw_cnn": dict(image=(3, 62, 47), convs=[ConvSpec(3, 18, 3, 1, 2), ConvSpec(18, 36, 3, 1, 2)], fc_in=5940,
                    n_classes=2),
}


def cnn_param_slices(arch: dict):
    """[(name, shape, start, end)] in named_parameters order."""
    out, off = [], 0
    for i, c in enumerate(arch["convs"]):
        for name, shape in ((f"conv{i}.weight", (c.cout, c.cin, c.k, c.k)), (f"conv{i}.bias", (c.cout,))):
            n = 1
            for s in shape:
                n *= s
            out.append((name, shape, off, off + n))
            off += n
    for name, shape in (("fc.weight", (arch["n_classes"], arch["fc_in"])), ("fc.bias", (arch["n_classes"],))):
        n = 1
        for s in shape:
            n *= s
        out.append((name, shape, off, off + n))
        off += n
    return out, off


def cnn_forward(arch: dict, params: dict, x: torch.Tensor) -> torch.Tensor:
    h = x
    for i, c in enumerate(arch["convs"]):
        h = F.conv2d(h, params[f"conv{i}.weight"], params[f"conv{i}.bias"], stride=1, padding=c.pad)
        h = F.relu(h)
        if c.pool > 1:
            h = F.max_pool2d(h, c.pool, c.pool)
    h = h.reshape(h.shape[0], -1)
    return F.linear(h, params["fc.weight"], params["fc.bias"])


def _unflatten(arch, w):
    slices, _ = cnn_param_slices(arch)
    return {name: w[a:b].view(shape) for name, shape, a, b in slices}


def cnn_local_step(arch: dict, w32: torch.Tensor, Xb: torch.Tensor, yb: torch.Tensor, clip: float, out: torch.Tensor):
    """Per-peer minibatch gradients through torch.func (library convolution kernels).  The hand-written
    sm_100a path is ops.cnn_native (used automatically on CUDA when available)."""
    from torch.func import functional_call, grad, vmap  # noqa: F401

    P, B = Xb.shape[0], Xb.shape[1]
    params = _unflatten(arch, w32.float())
    img = arch["image"]

    def loss_fn(p, x, y):
        logits = cnn_forward(arch, p, x.view(B, *img))
        return F.cross_entropy(logits, y.long())

    g = vmap(grad(loss_fn), in_dims=(None, 0, 0))(params, Xb.float(), yb)
    slices, d = cnn_param_slices(arch)
    flat = torch.cat([g[name].reshape(P, -1) for name, _, _, _ in slices], dim=1)
    total = flat.norm(dim=1, keepdim=True)
    coef = torch.clamp(clip / (total + 1e-6), max=1.0)
    out[:, :d] = -(flat * coef)
    if out.shape[1] > d:
        out[:, d:] = 0
    return out


class Model:
    def __init__(self, cfg: BiscottiConfig):
        self.cfg = cfg
        self.spec = DATASETS[cfg.dataset]
        self.kind = self.spec.model
        self.d = self.spec.n_params
        if self.kind in CNN_ARCH:
            self.arch = CNN_ARCH[self.kind]
            _, d = cnn_param_slices(self.arch)
            assert d == self.d, (d, self.d)

    def init_weights(self) -> torch.Tensor:
        """Genesis weights are zeros for linear models (block.go:46-53).  A zero-initialised CNN has zero
        gradients through ReLU, so CNNs start from a seeded small random vector instead (documented)."""
        if self.kind in CNN_ARCH:
            g = torch.Generator().manual_seed(self.cfg.seed * 31 + 7)
            return (0.01 * torch.randn(self.d, generator=g)).double()
        return torch.zeros(self.d, dtype=torch.float64)

    def local_step(self, w32, X, y, idx, out, peer_ids=None, slot=0):
        cfg = self.cfg
        if self.kind == "softmax":
            U, _ = ml.softmax_step(w32, X, y, idx, self.spec.n_classes, cfg.clip_norm, out=out, ldu=out.shape[1])
            return U
        if self.kind == "logistic":
            own = cfg.noise_scale if (cfg.noising and cfg.epsilon > 0) else 0.0
            return ml.logistic_step(w32, X, y, idx, cfg.logistic_alpha, cfg.logistic_lambda, own_noise_scale=own,
                                    seed=cfg.seed, peer_ids=peer_ids, slot=slot, out=out, ldu=out.shape[1])
        P = X.shape[0]
        B = idx.shape[1]
        Fd = X.shape[2]
        Xb = torch.gather(X, 1, idx.long().unsqueeze(-1).expand(P, B, Fd))
        yb = torch.gather(y, 1, idx.long())
        if X.is_cuda:
            from ..ops import cnn_native as cnn_ops

            if cnn_ops.supports(self.kind):
                return cnn_ops.local_step(self.kind, w32, Xb, yb, cfg.clip_norm, out)
        return cnn_local_step(self.arch, w32, Xb, yb, cfg.clip_norm, out)

    def eval_errors(self, Wc, base, X, y):
        if self.kind == "softmax":
            return ml.eval_softmax_errors(Wc, base, X, y, self.spec.n_classes)
        if self.kind == "logistic":
            return ml.eval_logistic_errors(Wc, base, X, y)
        errs = []
        for m in range(Wc.shape[0]):
            w = Wc[m, : self.d].float()
            if base is not None:
                w = w + base[: self.d].float()
            if X.is_cuda:
                from ..ops import cnn_native

                e = cnn_native.eval_errors(self.kind, w, X, y)
                if e is not None:
                    errs.append(e)
                    continue
            logits = cnn_forward(self.arch, _unflatten(self.arch, w), X.float().view(X.shape[0], *self.arch["image"]))
            errs.append((logits.argmax(-1) != y.long()).float().mean())
        return torch.stack(errs)
