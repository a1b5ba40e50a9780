"""Model branching — the reference's older centralised prototype (CentralBlockML/, SURVEY 2.5, optional capability).

A server keeps several *branches* of the model, each a chain of (weights, last aggregated gradient).  Every round
each client evaluates its private gradient on every branch, submits it to the branch whose previous gradient it
matches best, and every branch that received gradients appends their mean (CentralBlockML/code/modelBranch.py:4-37,
client.py:26-39).  Poisoners and honest clients thereby tend to separate onto different branches; a client
reports the branch with the lowest test error (client.py:46-56).

Quirk kept behind `metric`: the reference ranks branches by `scipy.spatial.distance.cosine` (= 1 - cosine
similarity) and keeps the *largest* value, i.e. it picks the least similar branch although the comment says "most
similar".  `metric="distance"` reproduces that, `metric="similarity"` (default) does what was meant.

DP noise: the prototype draws each client's noise from the density exp(-eps*||x||/2) with an emcee MCMC chain
(softmax_model_obj.py).  That density has a closed-form sampler — uniform direction, Gamma(d, 2/eps) radius — which
`sample_dp_noise` uses instead of a Markov chain.

All clients of a round are evaluated in one batched call per branch on the engine's kernels.
"""
from __future__ import annotations

import math

import torch

from ..config import BiscottiConfig
from ..data.synthetic import make_data
from ..models.registry import Model
from ..ops import ml


class ModelBranch:
    def __init__(self, w0: torch.Tensor, g0: torch.Tensor):
        self.chain = [(w0.clone(), g0.clone())]
        self.pending: list[torch.Tensor] = []
        self.history: list[int] = []

    @property
    def weights(self) -> torch.Tensor:
        return self.chain[-1][0]

    @property
    def previous_grad(self) -> torch.Tensor:
        return self.chain[-1][1]

    def submit(self, grad: torch.Tensor):
        self.pending.append(grad)

    def update(self) -> bool:
        """modelBranch.py:10-25: new gradient = mean of the pending ones, w += it."""
        if not self.pending:
            return False
        g = torch.stack(self.pending).mean(0)
        self.chain.append((self.weights + g, g))
        self.history.append(len(self.pending))
        self.pending = []
        return True


def sample_dp_noise(n: int, d: int, epsilon: float, generator=None, device="cpu") -> torch.Tensor:
    """n draws from the density proportional to exp(-epsilon * ||x|| / 2) on R^d."""
    if epsilon <= 0:
        return torch.zeros(n, d, device=device)
    direction = torch.randn(n, d, generator=generator, device=device)
    direction = direction / direction.norm(dim=1, keepdim=True).clamp_min(1e-30)
    # Gamma(shape d, scale 2/eps) as a sum of d exponentials is exact but O(d); use the library sampler
    radius = torch.distributions.Gamma(torch.tensor(float(d)), torch.tensor(epsilon / 2.0)).sample((n,)).to(device)
    return direction * radius[:, None]


def run_branching(dataset: str = "mnist", clients: int = 10, poisoners: int = 3, branches: int = 2, rounds: int = 30,
                  batch_size: int = 10, epsilon: float = 0.0, metric: str = "similarity", seed: int = 1, device: str = "cpu"):
    """CentralBlockML/code/main.py: returns per-round branch choices and the final per-branch test error / attack rate."""
    n = clients
    cfg = BiscottiConfig(num_nodes=n, dataset=dataset, poison_fraction=(poisoners / n if poisoners else 0.0),
                         batch_size=batch_size, seed=seed, commitments=False, secure_agg=False)
    dev = torch.device(device)
    data = make_data(cfg, list(range(n)), dev)
    model = Model(cfg)
    d = cfg.dim
    gen = torch.Generator(device=dev).manual_seed(seed)
    w0 = model.init_weights().float().to(dev)
    # branches start from the same weights with different random "previous gradients" (main.py seeds them randomly)
    bs = [ModelBranch(w0, 0.01 * torch.randn(d, generator=gen, device=dev)) for _ in range(branches)]
    U = torch.zeros(n, ml.pad_to(d, 4), dtype=torch.float32, device=dev)
    log = []
    for r in range(rounds):
        gen.manual_seed(seed * 7919 + r)
        idx = torch.rand(n, cfg.shard_size, generator=gen, device=dev).topk(batch_size, dim=1, largest=False).indices.int()
        grads, sims = [], []
        for b in bs:
            model.local_step(b.weights, data.X, data.y, idx, U, peer_ids=data.peer_ids, slot=r % 100)
            g = U[:, :d].clone()
            if epsilon > 0:
                g = g + sample_dp_noise(n, d, epsilon, gen, dev) / batch_size
            grads.append(g)
            sims.append(torch.nn.functional.cosine_similarity(g, b.previous_grad[None, :], dim=1))
        S = torch.stack(sims, dim=1)                                    # [clients, branches]
        score = (1.0 - S) if metric == "distance" else S                # reference quirk: largest cosine *distance*
        choice = score.argmax(dim=1)
        for c in range(n):
            bs[int(choice[c])].submit(grads[int(choice[c])][c])
        for b in bs:
            b.update()
        log.append(choice.tolist())
    out = []
    for b in bs:
        w = b.weights.unsqueeze(0)
        out.append({"blocks": len(b.chain) - 1,
                    "test_error": float(model.eval_errors(w, None, data.X_test, data.y_test)[0]),
                    "attack_rate": float(model.eval_errors(w, None, data.X_attack, data.y_attack)[0])})
    honest = n - poisoners
    last = log[-1]
    sep = {"honest_on_branch": [sum(1 for c in range(honest) if last[c] == k) for k in range(branches)],
           "poisoners_on_branch": [sum(1 for c in range(honest, n) if last[c] == k) for k in range(branches)]}
    return {"branches": out, "last_round": sep, "choices": log}
