"""Evaluation tooling, the go-python-compatible ML façade and the research extras (CPU)."""
import io

import numpy as np
import torch

from biscotti_b200 import ml_api
from biscotti_b200.config import BiscottiConfig
from biscotti_b200.data import utils as du
from biscotti_b200.eval import experiments as ex, parse_logs
from biscotti_b200.models import extras
from biscotti_b200.protocol.engine import BiscottiEngine


def test_ml_api_facade_round_trip():
    d = ml_api.init("creditcard", "creditcard3", epsilon=2.0, batch_size=10)
    assert d == 25
    w = np.zeros(d)
    e0 = ml_api.getTestErr(w)
    for _ in range(200):
        w = w + ml_api.privateFun(w)
    assert ml_api.getTestErr(w) < e0
    assert ml_api.getNoise(3).shape == (d,) and np.allclose(ml_api.getNoise(3), ml_api.getNoise(103))
    assert isinstance(ml_api.roni(w, np.zeros(d)), float) and abs(ml_api.roni(w, np.zeros(d))) < 1e-12
    deltas = np.random.RandomState(0).randn(10, d)
    deltas[2] += 20
    good = ml_api.krum(deltas, ml_api.default_clip(10))
    assert len(good) == 5 and 2 not in good
    assert 0.0 <= ml_api.get17AttackRate(w) <= 1.0


def test_log_markers_parse_like_the_reference_scraper():
    buf = io.StringIO()
    e = BiscottiEngine(BiscottiConfig(num_nodes=10, dataset="creditcard", max_iterations=3, shard_size=64, test_size=100),
                       log_stream=buf)
    e.run()
    res = parse_logs.parse(buf.getvalue().splitlines())
    assert [i for i, _ in res["train_error"]] == [0, 1, 2, 3]
    assert len(res["attack_rate"]) == 4 and res["s_per_iteration"] > 0


def test_cost_breakdown_and_closed_forms():
    rows = ex.cost_breakdown(peer_counts=(10,), dataset="creditcard", blocks=3, device="cpu", shard_size=64, test_size=100)
    r = rows[0]
    assert r["peers"] == 10 and r["total_ms"] > 0
    assert abs(r["total_ms"] - (r["noising_ms"] + r["verification_ms"] + r["aggregation_ms"] + r["flooding_ms"] + r["local_ms"])) < 1e-6
    cap = ex.committee_capture(100, stakes=(0.1, 0.3), thresholds=(0.01,))
    assert cap[0]["committee_size"] < cap[1]["committee_size"]
    um = ex.unmasked_probability(stakes=(0.3,), noisers=(2, 5))
    assert um[0]["p_unmasked"] > um[1]["p_unmasked"] > 0


def test_churn_rejoin_and_data_utils_and_extras():
    rows = ex.churn(period_blocks=(2,), peers=10, dataset="creditcard", blocks=6, device="cpu")
    assert rows[0]["rejoin_ok"]
    X = np.random.RandomState(1).randn(100, 5)
    y = (np.arange(100) % 10).astype(float)
    Xs, mu, sg = du.standardize_cols(X)
    assert np.allclose(Xs.mean(0), 0, atol=1e-9) and np.allclose(Xs.std(0), 1)
    shards = du.slice_uniform(X, y, 4)
    assert len(shards) == 4 and shards[0].shape == (25, 6)
    bad = du.generate_poisoned(X, y)
    assert not (bad[:, -1] == 1).any() and (bad[:, -1] == 7).sum() == 20
    deltas = torch.randn(8, 6)
    deltas[5] += 10
    assert extras.single_krum(deltas, 2) != 5
    assert extras.pick_branch_by_cosine(torch.ones(4), [-torch.ones(4), torch.ones(4)]) == 1
    d, h = extras.linear_topk_step(torch.zeros(5), torch.randn(20, 5), torch.randn(20), torch.zeros(5), topk=2)
    assert int((d != 0).sum()) == 2
    g = extras.svm_step(torch.zeros(4 * 6 + 4), torch.randn(10, 6), torch.randint(0, 4, (10,)), 4)
    assert g.shape == (28,)
