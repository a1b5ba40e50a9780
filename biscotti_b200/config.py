"""Run configuration: every flag and compile-time constant of the reference peer, in one place.

Flag parity: DistSys/main.go:613-647 (``-t -i -d -f -a -pa -p -c -na -nv -nn -sa -np -vp -ep -po -ns
-rs``), constants DistSys/main.go:28-60 and DistSys/honest.go:44-57, dataset table
ML/Pytorch/datasets.py:18-52 and DistSys/localTest.sh:16-20.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field, asdict

BASE_PORT = 8000           # main.go:29 (kept for the peers-file format only; no sockets are opened)
DEFAULT_STAKE = 10         # main.go:39
STAKE_UNIT = 5             # honest.go:46
VERIFIER_PRIME, MINER_PRIME, NOISER_PRIME = 2, 3, 5   # main.go:41-43
PRECISION = 4              # main.go:45
POLY_SIZE = 10             # main.go:46
MAX_ITERATIONS = 100       # main.go:48
FAIL_PROB = -0.005         # main.go:55 (fault injection off)
CONV_THRESHOLD = 0.0       # honest.go:51
NOISE_TABLE_ITERS = 100    # client_obj.py:17, logistic_model.py:21
# The reference's phase timeouts (main.go:31-36, 786-821) have no counterpart: arrival orders are explicit
# seeded permutations and a missed deadline is the empty-block branch of the engine, not a wall-clock wait.


@dataclass(frozen=True)
class DatasetSpec:
    name: str
    model: str          # softmax | logistic | mnist_cnn | cifar_cnn | lfw_cnn | svm
    n_features: int     # flattened input size
    n_classes: int
    n_params: int
    image_shape: tuple = ()


# ML/Pytorch/datasets.py + model files (param counts: SURVEY §2.5)
DATASETS = {
    "mnist": DatasetSpec("mnist", "softmax", 784, 10, 7850, (1, 28, 28)),
    "creditcard": DatasetSpec("creditcard", "logistic", 25, 2, 25),
    "cifar": DatasetSpec("cifar", "cifar_cnn", 3072, 10, 259770, (3, 32, 32)),
    "mnist_cnn": DatasetSpec("mnist_cnn", "mnist_cnn", 784, 10, 164266, (1, 28, 28)),
    "lfw": DatasetSpec("lfw", "lfw_cnn", 8742, 2, 18254, (3, 62, 47)),
    "lfw_softmax": DatasetSpec("lfw_softmax", "softmax", 8742, 2, 17486, (3, 62, 47)),
}


@dataclass
class BiscottiConfig:
    # ---- reference flags --------------------------------------------------------------------
    num_nodes: int = 10               # -t
    dataset: str = "creditcard"       # -d
    num_miners: int = 3               # -na
    num_verifiers: int = 3            # -nv
    num_noisers: int = 2              # -nn
    secure_agg: bool = True           # -sa
    noising: bool = True              # -np
    verification: bool = True         # -vp
    epsilon: float = 2.0              # -ep
    poison_fraction: float = 0.0      # -po
    perc_samples: int = 70            # -ns (percent of peers whose updates Krum waits for)
    random_sampling: bool = False     # -rs
    colluders: int = 0                # -c
    defense: str = "KRUM"             # POISON_DEFENSE: KRUM | RONI
    # ---- constants exposed as knobs -----------------------------------------------------------
    max_iterations: int = MAX_ITERATIONS
    batch_size: int = 10              # honest.go:47
    precision: int = PRECISION
    fail_prob: float = FAIL_PROB
    partitions: tuple = ()            # ((peer, first_block, last_block), ...): peer is unreachable for those blocks
                                      # (blockNode.sh: iptables DROP on the peer's port for 30 s, then flush)
    krum_adversary_frac: float = 0.5  # main.go:780-783
    clip_norm: float = 100.0          # client.py:53
    dp_delta: float = 1e-5            # client_obj.py:61
    logistic_alpha: float = 1e-2      # logistic_model.py
    logistic_lambda: float = 1e-2
    dp_in_model: bool = False         # main.go:155 DP_IN_MODEL
    roni_threshold: float = 0.02      # honest.go:598-629
    # ---- framework knobs (no reference counterpart) -------------------------------------------
    mode: str = "biscotti"            # biscotti | fedsys
    seed: int = 1
    srs_tau: int = 2                  # publicKey.go:36 compat; 0 => random tau
    commitments: bool = True          # compute KZG commitments + witnesses (reference always does)
    sign_approvals: bool = True       # verifiers Schnorr-sign accepted commitments (kyber.go:873-896)
    share_trigger: str = "half"       # "half": miners act on NUM_SAMPLES/2 shares (main.go:360); "all"
    deterministic_time: bool = True   # inject block timestamps so chains are reproducible
    shard_size: int = 600             # rows per peer shard (mnist 60000/100)
    test_size: int = 2000
    eval_every: int = 1               # reference evaluates every block (honest.go:141-162)
    msm_window: int | None = None     # fixed-base window bits for the MSM tables (None = largest that fits HBM budget)
    msm_table_gb: float | None = None # HBM budget of the MSM window tables in GiB (None = half of the device; the
                                      # BISCOTTI_MSM_TABLE_GB environment variable overrides both).  The constructor
                                      # picks the widest table layout that fits: two-width (13,8) / (12,9) / (11,10) bit
                                      # windows, then uniform 7..3 bit windows; 0 = no tables: every MSM runs window-free
                                      # (Straus over the bases themselves, `msm_groups_kernel`), memory = the bases only
    affine_wire: bool = False         # True: normalise every chunk commitment / witness to the 64-byte affine
                                      # wire form (kyber.go:88-168) even though no wire exists between GPUs
    data_sep: float = 3.6             # synthetic data: distance scale of the class prototypes.  3.6 makes the ten classes
                                      # overlap like MNIST does for a linear model: test error passes 0.10 near block 80 and
                                      # ends at ~0.09 after 102 blocks (reference: 0.095), and - what matters for timing - the
                                      # updates keep realistic magnitudes.  With well separated classes (6.0, round 1) the
                                      # error is 0.002 after a few blocks and most coordinates of an update quantise to
                                      # zero, which lets the commitment / witness MSMs skip most of their work
    data_noise: float = 1.0           # synthetic data ("blobs"): per-feature Gaussian noise around the prototype
    data_model: str = "blobs"         # "blobs": prototype + isotropic noise of std data_noise in every feature.  Its within-class
                                      # variation is 784-dimensional: learning is slow (the default above) and updates keep
                                      # non-zero coordinates everywhere - the demanding case for the MSMs, hence the default
                                      # for timing - but every update is about equally far from every other, so under the
                                      # reference's DP noise (std 0.54 per coordinate at epsilon 2, norm 48) Multi-Krum cannot
                                      # tell poisoners from honest peers.
                                      # "lowrank": class c = data_radius * (axis of its pair c // 2) + its own axis (two classes
                                      # of a pair sit data_gap apart), spread data_spread along the class axes, data_latent
                                      # class-neutral directions of std data_latent_sigma, per-pixel noise data_iso.  Rows of one
                                      # class look alike, as digit images do, so peers that all train on one relabelled class
                                      # send the tight, far-off cluster of updates Multi-Krum is built to reject (squared
                                      # distance to an honest update ~1700 against a noise fluctuation of ~40 per pair)
    data_radius: float = 0.0
    data_gap: float = 28.28
    data_spread: float = 6.0
    data_latent: int = 24
    data_latent_sigma: float = 2.5
    data_iso: float = 0.15
    data_dir: str | None = None       # directory of <dataset><i>.npy shards (data/files.py); None = synthetic data
    prefetch: bool = True             # queue iteration t+1's local step + commitment MSM behind iteration t's
                                      # aggregation kernel so the device is busy while the host assembles block t
    ring_depth: int = 3               # slots of the ring of per-block device buffers / events / command lists (>= 2)
    pipeline: bool = True             # GPU: the block's dependency chain on a high-priority stream, everything off the
                                      # chain (witness MSM, miners' aggregation, error rates, next local step +
                                      # commitments) on a second stream, the approvals on a third.  False = one stream,
                                      # strictly in program order (what an instrumented run uses)
    lookahead: bool = True            # pipelined GPU engines: plan block t+1 and queue the front of its chain (VRF ->
                                      # noisers -> noise -> Krum -> plan -> share sums; scratch writes only) from inside
                                      # step t, right after block t is hashed, so the device never waits for the host to
                                      # come back around the loop
    native_schedule: bool = True      # pipelined GPU engines, standard configuration: replay the block's launches as native
                                      # command lists (protocol/schedule.py) instead of issuing them one by one from Python
    gram_impl: str = "auto"           # auto | tcgen05 | simt
    comm_impl: str = "fused"          # fused (in-kernel P2P) | nccl (baseline)
    extra: dict = field(default_factory=dict)

    # ---- derived quantities (main.go:672-684, 825) ---------------------------------------------
    @property
    def spec(self) -> DatasetSpec:
        return DATASETS[self.dataset]

    @property
    def dim(self) -> int:
        return self.spec.n_params

    @property
    def num_samples(self) -> int:
        n = int(self.num_nodes * self.perc_samples / 100.0)
        cap = self.num_nodes - self.num_verifiers - self.num_miners
        return max(1, min(n, cap))

    @property
    def krum_thresh(self) -> int:
        if self.random_sampling:
            return max(1, self.num_nodes - self.num_verifiers - self.num_miners)
        return self.num_samples

    @property
    def total_shares(self) -> int:
        return int(math.ceil(2.0 * POLY_SIZE / self.num_miners)) * self.num_miners

    @property
    def shares_per_miner(self) -> int:
        return self.total_shares // self.num_miners

    @property
    def nchunk(self) -> int:
        return (self.dim + POLY_SIZE - 1) // POLY_SIZE

    @property
    def noise_sigma(self) -> float:
        if self.epsilon <= 0:
            return 0.0
        if self.spec.model == "logistic":
            return math.sqrt(2.0 * math.log(1.25)) / self.epsilon           # logistic_model.py:81
        return math.sqrt(2.0 * math.log(1.25 / self.dp_delta)) / self.epsilon  # client_obj.py:61-63

    @property
    def noise_scale(self) -> float:
        """getNoise scale incl. the sum over `batch` presampled normals (std sqrt(batch)*sigma)."""
        s = self.noise_sigma * math.sqrt(self.batch_size)
        if self.spec.model == "logistic":
            return -self.logistic_alpha / self.batch_size * s                # logistic_model.py:108-109
        return -1.0 / self.batch_size * s                                     # client_obj.py:97-98

    def to_dict(self) -> dict:
        return asdict(self)
