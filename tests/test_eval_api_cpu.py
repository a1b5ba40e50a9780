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
    assert bad.shape == (10, 6) and (bad[:, -1] == 7).all()                    # the ten source rows, all labelled 7
    deltas = torch.randn(8, 6)
    deltas[5] += 10
    assert extras.single_krum(deltas, 2) != 5
    assert extras.pick_branch_by_cosine(torch.ones(4), [-torch.ones(4), torch.ones(4)]) == 1
    d, h = extras.linear_topk_step(torch.zeros(5), torch.randn(20, 5), torch.randn(20), torch.zeros(5), topk=2)
    assert int((d != 0).sum()) == 2
    g = extras.svm_step(torch.zeros(4 * 6 + 4), torch.randn(10, 6), torch.randint(0, 4, (10,)), 4)
    assert g.shape == (28,)


def test_file_backed_shards_round_trip(tmp_path):
    """make-shards layout (parse_mnist.py) -> load_peer_data -> an engine run from files."""
    import numpy as np

    from biscotti_b200.config import BiscottiConfig
    from biscotti_b200.data.files import load_peer_data, load_rows, write_shards
    from biscotti_b200.protocol.engine import BiscottiEngine

    rs = np.random.RandomState(0)
    n, F = 400, 784
    y = rs.randint(0, 10, size=n).astype(np.float64)
    X = rs.randn(n, F) + y[:, None] * 0.3
    files = write_shards(str(tmp_path), "mnist", X, y, X[:100], y[:100], n_shards=4)
    names = {f.split("/")[-1] for f in files}
    assert {"mnist0.npy", "mnist3.npy", "mnist_train.npy", "mnist_test.npy", "mnist_digit1.npy", "mnist_bad.npy"} <= names
    Xb, yb = load_rows(str(tmp_path / "mnist_bad"), is_train=False, train_cut=0.0)
    assert len(yb) == int((y == 1).sum()) and (yb == 7).all()                  # the digit-1 rows, all labelled 7
    Xtr, ytr = load_rows(str(tmp_path / "mnist0"))
    assert Xtr.shape == (80, F)                                               # first 80 % of a 100-row shard
    cfg = BiscottiConfig(num_nodes=4, dataset="mnist", shard_size=80, poison_fraction=0.5, data_dir=str(tmp_path),
                         max_iterations=2, commitments=False, num_verifiers=1, num_miners=1, num_noisers=1)
    d = load_peer_data(cfg, [0, 1, 2, 3], str(tmp_path))
    assert d.X.shape == (4, 80, F) and bool(d.poisoned[3]) and not bool(d.poisoned[0])
    assert not (d.y[3] == 1).any()                                            # the poisoner reads mnist_bad
    eng = BiscottiEngine(cfg)
    eng.run()
    assert len(eng.chain) == 4 and eng.chain.verify()


def test_ml_sandboxes_and_lsh_sieve():
    import torch

    from biscotti_b200.sim import ml_main

    rows = ml_main.run_clients("creditcard", honest=6, poisoners=0, rounds=12, defense="none", log_every=11)
    assert rows[-1]["test_error"] < rows[0]["test_error"] + 0.05 and rows[-1]["accepted"] == 6
    # the ML-only aggregator of client.py:88-108 (mean gradient, SGD with momentum and weight decay)
    sg = ml_main.run_clients("mnist", honest=6, poisoners=0, rounds=30, optimizer="sgd", lr=0.01, log_every=29, data_sep=6.0)
    assert sg[-1]["test_error"] < 0.02 < sg[0]["test_error"]
    kr = ml_main.run_clients("mnist", honest=6, poisoners=4, rounds=3, defense="krum", log_every=1)
    assert all(r["accepted"] == 6 for r in kr)                                # n - clip survive Multi-Krum
    leak = ml_main.gradient_leak("mnist", batches=(1, 10), epsilons=(0.0,))
    assert leak[0]["leak_cosine"] > 0.999 and leak[1]["leak_cosine"] < leak[0]["leak_cosine"]
    g = torch.Generator().manual_seed(0)
    base = torch.randn(1, 64, generator=g)
    deltas = torch.cat([base + 0.05 * torch.randn(8, 64, generator=g), -base + 0.05 * torch.randn(2, 64, generator=g)])
    keep = ml_main.lsh_sieve(deltas, keep_fraction=0.8)
    assert set(keep.tolist()) == set(range(8))                                # the two outliers are sieved out


def test_gradient_check_and_noise_utils():
    import numpy as np

    from biscotti_b200.data.utils import check_gradient, exp_noise, laplace_noise

    A = np.random.RandomState(1).randn(5, 5)
    A = A @ A.T

    def quad(w):
        return 0.5 * float(w @ A @ w), A @ w

    assert check_gradient(quad, np.arange(5.0)) < 1e-5
    try:
        check_gradient(lambda w: (float(w @ A @ w), A @ w), np.arange(5.0))   # wrong by a factor 2
        raise RuntimeError("should have failed")
    except AssertionError:
        pass
    assert abs(np.mean(np.abs(laplace_noise(20000, 2.0))) - 2.0) < 0.1
    assert abs(np.mean(exp_noise(20000, 3.0)) - 3.0) < 0.15


def test_central_block_model_branching():
    """CentralBlockML: branches advance by the mean of the gradients submitted to them; the DP-noise sampler has
    the radius distribution of exp(-eps ||x|| / 2)."""
    import torch

    from biscotti_b200.sim import central_block as cb

    b = cb.ModelBranch(torch.zeros(3), torch.ones(3))
    assert not b.update()
    b.submit(torch.tensor([1.0, 0.0, 0.0]))
    b.submit(torch.tensor([3.0, 2.0, 0.0]))
    assert b.update() and torch.allclose(b.weights, torch.tensor([2.0, 1.0, 0.0])) and b.history == [2]
    assert torch.allclose(b.previous_grad, torch.tensor([2.0, 1.0, 0.0]))
    g = torch.Generator().manual_seed(0)
    z = cb.sample_dp_noise(4000, 8, 2.0, g)
    assert abs(float(z.norm(dim=1).mean()) - 8 * 2 / 2.0) < 0.3            # E||x|| = d * 2/eps
    res = cb.run_branching("creditcard", clients=8, poisoners=2, branches=2, rounds=10)
    assert len(res["branches"]) == 2 and len(res["choices"]) == 10
    assert sum(res["last_round"]["honest_on_branch"]) == 6 and sum(res["last_round"]["poisoners_on_branch"]) == 2
    assert sum(x["blocks"] for x in res["branches"]) >= 10
    quirk = cb.run_branching("creditcard", clients=8, poisoners=2, branches=2, rounds=3, metric="distance")
    assert quirk["choices"] != cb.run_branching("creditcard", clients=8, poisoners=2, branches=2, rounds=3)["choices"]


def test_chrome_trace_of_an_instrumented_run(tmp_path):
    """localtest --trace: one duration event per phase per block plus one per block, valid JSON for chrome://tracing."""
    import json
    import subprocess
    import sys
    import os

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = tmp_path / "trace.json"
    r = subprocess.run([sys.executable, "-m", "biscotti_b200", "localtest", "10", "creditcard", "--iterations", "3",
                        "--device", "cpu", "--shard-size", "100", "--no-dump", "--trace", str(out)],
                       capture_output=True, text=True, cwd=root, env=dict(os.environ, PYTHONPATH=root), timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    tr = json.load(open(out))
    ev = [e for e in tr["traceEvents"] if e.get("ph") == "X"]
    blocks = [e for e in ev if e["cat"] == "block"]
    phases = [e for e in ev if e["cat"] == "phase"]
    assert len(blocks) == 4 and {"local_step", "noise", "verify", "block"} <= {e["name"] for e in phases}
    assert all(e["dur"] >= 0 and e["ts"] >= 0 for e in ev)
    starts = [e["ts"] for e in blocks]
    assert starts == sorted(starts)
