"""Protocol engine on the GPU: MNIST-shaped softmax peers, determinism, learning, poisoning defence."""
import pytest
import torch

from biscotti_b200.config import BiscottiConfig
from biscotti_b200.protocol.engine import BiscottiEngine

pytestmark = pytest.mark.gpu


def _cfg(**kw):
    base = dict(num_nodes=30, dataset="mnist", max_iterations=4, shard_size=100, test_size=500, msm_window=7, data_sep=6.0)
    base.update(kw)
    return BiscottiConfig(**base)


def test_mnist_chain_on_gpu_is_deterministic_and_consistent(lib):
    lib.bsc_launch_count_reset()
    a = BiscottiEngine(_cfg(), device="cuda:0")
    b = BiscottiEngine(_cfg(), device="cuda:0")
    ra, rb = a.run(), b.run()
    assert lib.bsc_launch_count() > 50
    assert a.chain == b.chain and a.chain.verify()
    assert all(r.bad_shares == 0 for r in ra)
    assert any(not r.empty for r in ra)
    assert a.print_chain() == b.print_chain()


def test_mnist_learns_and_krum_rejects_poisoners():
    # rows of a class look alike ("lowrank"), poisoners train on source-class rows relabelled as the target class
    # (mnist_bad), DP noise light enough (epsilon 8) for the distances to show: Multi-Krum must keep every poisoner out
    cfg = _cfg(num_nodes=40, max_iterations=14, poison_fraction=0.3, epsilon=8.0, data_model="lowrank")
    e = BiscottiEngine(cfg, device="cuda:0")
    reps = e.run()
    assert len(e.poisoners) == 11 and e.chain.verify()
    assert all(r.n_contributors >= 2 for r in reps)
    assert not any(c in e.poisoners for r in reps for c in r.contributors)
    assert reps[-1].test_error < 0.12 and max(r.attack_rate for r in reps[5:]) < 0.3


def test_fedsys_mode_on_gpu():
    cfg = _cfg(mode="fedsys", perc_samples=35, max_iterations=6)
    e = BiscottiEngine(cfg, device="cuda:0")
    reps = e.run()
    assert reps[-1].test_error < reps[0].test_error


def test_table_free_msm_mode_gives_the_same_chain():
    """msm_table_gb = 0: no window tables, every commitment / witness MSM runs window-free over the bases themselves
    (msm_groups_kernel) - the mode for models whose tables would not fit.  Same group elements, same ledger."""
    a = BiscottiEngine(_cfg(), device="cuda:0")
    b = BiscottiEngine(_cfg(msm_window=None, msm_table_gb=0.0), device="cuda:0")
    assert a.commit.c_table is not None and b.commit.c_table is None and b.commit.w_table is None
    a.run()
    b.run()
    assert a.chain == b.chain and b.chain.verify()
    a.close()
    b.close()
