"""Experiment drivers mirroring the reference's evaluation campaigns, run in-process on the engine.

Each function corresponds to one of the reference's `runEval.sh` + `parseLogs.py` pipelines (SURVEY
§2.6) and returns plain dict/list results (optionally written as CSV with the same column meaning):

  cost_breakdown        eval/eval_performance (40/60/80/100 peers: total, noising, verification,
                        aggregation, flooding) — perf_breakdown_vsFedSys.sh, parseLogs.py:79-276
  fedsys_scale          eval/eval_FedSys_scale (Biscotti vs FedSys time per iteration)
  committee_scale       eval/eval_vrf_scale (grow noisers / verifiers / aggregators: 3, 5, 10, 26)
  noise_krum            eval/eval_noise_krum (-po=0.3 -ns=70 -ep in {0.01 .. 5})
  poison_heatmap        eval/eval_poison_nsamples (-po x -ns heat-map of the 1->7 attack rate)
  privacy_utility       eval/eval_privacy_utility_krum (-np on/off, epsilon sweep)
  churn                 eval/eval_FT + DistSys/failAndRestartLocal.sh (kill/restart a random peer
                        every `period` blocks; restarted peers adopt the longest chain)
  committee_capture     eval/eval_vrf_security/vrf_security.py (closed-form binomial)
  unmasked_probability  eval/eval_privacy_noise_attack/vrf_noise_security.py (closed-form)

Phase names map to the reference's log scraping windows: noising = "Getting noise from" ->
"Sending update to verifiers"; verification = -> "Sending update to miners"; aggregation = share
generation + miner fold + reconstruction; flooding = block assembly/append (the remainder).
"""
from __future__ import annotations

import csv
import math
import time

from ..config import BiscottiConfig


def _engine(cfg: BiscottiConfig, device):
    from ..protocol.engine import BiscottiEngine

    return BiscottiEngine(cfg, device=device)


def _run(cfg: BiscottiConfig, device, blocks: int, timing: bool = True):
    if timing:
        # per-phase attribution: keep every phase inside its own iteration (the prefetch would book the next
        # iteration's local step + commitment MSM under this iteration's "block" phase)
        cfg.prefetch = False
    eng = _engine(cfg, device)
    eng._timing = timing
    t0 = time.perf_counter()
    reps = eng.run(blocks)
    dt = time.perf_counter() - t0
    eng.flush_reports()
    eng.close()      # frees the window tables (tens of GB): the next configuration of a campaign builds its own
    return None, reps, dt


def _phase_means(reps):
    keys = ["local_step", "commit", "noise", "verify", "witness", "miner_aggregate", "secagg", "block"]
    n = max(1, len(reps))
    m = {k: sum(r.phase_ms.get(k, 0.0) for r in reps) / n for k in keys}
    return {
        "noising_ms": m["noise"],
        "verification_ms": m["verify"],
        "aggregation_ms": m["commit"] + m["witness"] + m["miner_aggregate"] + m["secagg"],
        "flooding_ms": m["block"],
        "local_ms": m["local_step"],
        "total_ms": sum(m.values()),
    }


def cost_breakdown(peer_counts=(40, 60, 80, 100), dataset="mnist", blocks=20, device="cuda:0", **kw):
    rows = []
    for n in peer_counts:
        cfg = BiscottiConfig(num_nodes=n, dataset=dataset, max_iterations=blocks, **kw)
        eng, reps, dt = _run(cfg, device, blocks)
        row = {"peers": n, "s_per_block": dt / blocks, "final_error": reps[-1].test_error}
        row.update(_phase_means(reps[2:] or reps))
        rows.append(row)
    return rows


def fedsys_scale(peer_counts=(100, 200), dataset="mnist", blocks=20, device="cuda:0"):
    rows = []
    for n in peer_counts:
        for mode, ns in (("biscotti", 70), ("fedsys", 35)):
            cfg = BiscottiConfig(num_nodes=n, dataset=dataset, max_iterations=blocks, mode=mode, perc_samples=ns)
            eng, reps, dt = _run(cfg, device, blocks, timing=False)
            rows.append({"peers": n, "mode": mode, "s_per_block": dt / blocks, "final_error": reps[-1].test_error})
    return rows


def committee_scale(sizes=(3, 5, 10, 26), peers=100, dataset="mnist", blocks=10, device="cuda:0"):
    rows = []
    for size in sizes:
        for role in ("noisers", "verifiers", "aggregators"):
            kw = dict(num_noisers=2, num_verifiers=3, num_miners=3)
            kw[{"noisers": "num_noisers", "verifiers": "num_verifiers", "aggregators": "num_miners"}[role]] = size
            cfg = BiscottiConfig(num_nodes=peers, dataset=dataset, max_iterations=blocks, **kw)
            eng, reps, dt = _run(cfg, device, blocks, timing=False)
            rows.append({"committee": role, "size": size, "s_per_block": dt / blocks,
                         "total_shares": cfg.total_shares, "final_error": reps[-1].test_error})
    return rows


def noise_krum(epsilons=(0.01, 0.1, 0.5, 1.0, 2.0, 5.0), peers=100, dataset="mnist", blocks=100, poison=0.3,
               device="cuda:0"):
    rows = []
    for ep in epsilons:
        cfg = BiscottiConfig(num_nodes=peers, dataset=dataset, max_iterations=blocks, poison_fraction=poison, epsilon=ep)
        eng, reps, dt = _run(cfg, device, blocks, timing=False)
        tail = reps[int(0.8 * len(reps)):]
        rows.append({"epsilon": ep, "final_error": reps[-1].test_error,
                     "max_attack_rate_after_80pct": max(r.attack_rate or 0.0 for r in tail)})
    return rows


def poison_heatmap(fractions=(0.5, 0.45, 0.4, 0.3, 0.1), collected=(20, 40, 70), peers=100, dataset="mnist", blocks=100,
                   device="cuda:0", **kw):
    """plotHeatMap.py:46-88: max attack rate after block 80 for -po x -ns with -rs=true."""
    rows = []
    for po in fractions:
        for ns in collected:
            cfg = BiscottiConfig(num_nodes=peers, dataset=dataset, max_iterations=blocks, poison_fraction=po,
                                 perc_samples=ns, random_sampling=True, **kw)      # epsilon stays at the -ep default (2.0)
            eng, reps, dt = _run(cfg, device, blocks, timing=False)
            tail = reps[int(0.8 * len(reps)):]
            rows.append({"poisoners": po, "collected": ns, "max_attack_rate": max(r.attack_rate or 0.0 for r in tail)})
    return rows


def privacy_utility(epsilons=(0.1, 0.5, 1.0, 2.0), peers=100, dataset="mnist", blocks=100, device="cuda:0"):
    rows = []
    for noising in (True, False):
        for ep in epsilons if noising else (0.0,):
            cfg = BiscottiConfig(num_nodes=peers, dataset=dataset, max_iterations=blocks, noising=noising,
                                 epsilon=ep if noising else 2.0)
            eng, reps, dt = _run(cfg, device, blocks, timing=False)
            rows.append({"noising": noising, "epsilon": ep, "final_error": reps[-1].test_error})
    return rows


def churn(period_blocks=(16, 8, 4, 2), peers=50, dataset="mnist", blocks=100, device="cuda:0"):
    """Kill a random peer (never node 0) every `period` blocks and restart it `period` blocks later; on
    restart it re-bootstraps from the serialised chain and must end on the common chain
    (failAndRestartLocal.sh, main.go:1000-1013)."""
    import numpy as np

    from ..protocol.engine import BiscottiEngine

    rows = []
    for period in period_blocks:
        cfg = BiscottiConfig(num_nodes=peers, dataset=dataset, max_iterations=blocks)
        eng = BiscottiEngine(cfg, device=device)
        rs = np.random.RandomState(period)
        down = {}
        rejoin_ok = True
        t0 = time.perf_counter()
        for it in range(blocks):
            for p, back in list(down.items()):
                if it >= back:
                    eng.alive[p] = True
                    del down[p]
                    # the restarted peer receives the chain and must be able to adopt and verify it
                    probe = BiscottiEngine(BiscottiConfig(num_nodes=peers, dataset=dataset, commitments=False,
                                                          max_iterations=1, shard_size=8, test_size=8), device="cpu")
                    rejoin_ok = rejoin_ok and probe.adopt_chain(eng.chain.serialize())
            if it % period == 0 and it > 0:
                p = int(rs.randint(1, peers))
                eng.alive[p] = False
                down[p] = it + period
            eng.step()
        dt = time.perf_counter() - t0
        rows.append({"kill_period_blocks": period, "s_per_block": dt / blocks, "final_error": eng.reports[-1].test_error,
                     "empty_blocks": sum(r.empty for r in eng.reports), "rejoin_ok": rejoin_ok})
        eng.close()
    return rows


def time_to_accuracy(target=0.10, peers=100, dataset="mnist", blocks=102, device="cuda:0", **kw):
    """The reference's headline learning metric (eval/eval_performance/parseLogs.py:197-210): run MAX_ITERATIONS blocks
    (102 logged "Train Error" lines), take the wall-clock time of each block's log line, report seconds until the error
    first drops to `target` and seconds per block.  The engine runs its production schedule (pipelined, not instrumented);
    a block's time stamp is taken when step() returns, i.e. when the block is on the ledger."""
    cfg = BiscottiConfig(num_nodes=peers, dataset=dataset, max_iterations=blocks, **kw)
    eng = _engine(cfg, device)
    stamps = []
    t0 = time.perf_counter()
    for _ in range(blocks):
        eng.step()
        stamps.append(time.perf_counter() - t0)
    eng.flush_reports()
    errs = [r.test_error for r in eng.reports]
    hit = next((i for i, e in enumerate(errs) if e is not None and e <= target), None)
    tail = eng.reports[int(0.8 * blocks):]
    # the first two blocks carry one-off costs (buffer allocation, command-list construction): rate over the rest
    rate = (stamps[-1] - stamps[1]) / max(1, blocks - 2)
    out = {"peers": peers, "dataset": dataset, "blocks": blocks, "s_per_block": rate, "first_two_blocks_s": stamps[1],
           "target_error": target, "blocks_to_target": hit, "seconds_to_target": None if hit is None else stamps[hit],
           "final_error": errs[-1], "error_every_10": [round(e, 4) for e in errs[::10]],
           "max_attack_rate_after_80pct": max((r.attack_rate or 0.0) for r in tail),
           "empty_blocks": sum(r.empty for r in eng.reports), "chain_ok": eng.chain.verify()}
    eng.close()
    return out


# ---- closed-form committee analyses (pure Python in the reference as well) ---------------------------
def committee_capture(num_clients=100, stakes=(0.05, 0.1, 0.2, 0.3), thresholds=(0.01, 0.001)):
    """Smallest committee size for which P[adversary holds a majority] < threshold, binomial model of
    vrf_security.py (sum_{i > k/2} C(k, i) s^i (1-s)^(k-i))."""
    out = []
    for thr in thresholds:
        for s in stakes:
            size = None
            for k in range(3, num_clients):
                p = sum(math.comb(k, i) * s ** i * (1 - s) ** (k - i) for i in range(k // 2 + 1, k + 1))
                if p < thr:
                    size = k
                    break
            out.append({"threshold": thr, "adversary_stake": s, "committee_size": size})
    return out


def unmasked_probability(stakes=(0.05, 0.1, 0.2, 0.3), noisers=(2, 3, 5, 10), verifiers=3):
    """P[an update is unmasked] = P[all its noisers collude] * P[at least one verifier colludes]
    (vrf_noise_security.py; what the -c accounting of main.go:1026-1057 measures empirically)."""
    out = []
    for s in stakes:
        for nn in noisers:
            out.append({"adversary_stake": s, "noisers": nn, "p_unmasked": (s ** nn) * (1 - (1 - s) ** verifiers)})
    return out


def write_csv(rows, path):
    if not rows:
        return
    with open(path, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(rows[0].keys()))
        w.writeheader()
        w.writerows(rows)
