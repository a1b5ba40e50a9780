#!/usr/bin/env python
"""Full-length campaigns on one GPU (results -> gpurun_out/campaigns.json, copied to profiles/ by hand):

  * time to accuracy, MNIST-shaped softmax, 100 peers, 102 blocks, reference defaults (parseLogs.py:197-210)
  * the same with 30 % poisoners and epsilon = 0.5 (BASELINE.json config 3)
  * poisoning heat-map rows with -rs=true (plotHeatMap.py:46-88)
  * CIFAR-shaped CNN, 200 peers, 50 blocks on the ledger path (raw gradient sums, no learning rate - honest.go:358-375),
    and the same CNN kernels under the reference's ML-only aggregator (mean gradient + SGD, client.py:30, 88-108)

The synthetic classes overlap (`data_sep` = 3.6, the package default) so that the error needs tens of blocks to pass 0.10.
The poisoning runs are done on both synthetic data models (config.py: data_model): "blobs", where every update is about
equally far from every other and Multi-Krum has nothing to go on under the DP noise, and "lowrank", where rows of a class
look alike and the poisoners' updates form a far-off cluster."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.eval import experiments as ex  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--sep", type=float, nargs="*", default=[3.6])
ap.add_argument("--what", nargs="*", default=["tta", "poison", "heatmap", "cifar", "cifar_sgd"])
ap.add_argument("--blocks", type=int, default=102)
ap.add_argument("--models", nargs="*", default=["blobs", "lowrank"])
a = ap.parse_args()
out = {}
t0 = time.time()
for sep in a.sep:
    kw = dict(data_sep=sep)
    if "tta" in a.what:
        out[f"tta_sep{sep}"] = ex.time_to_accuracy(0.10, blocks=a.blocks, **kw)
        print("TTA", sep, json.dumps(out[f"tta_sep{sep}"]), flush=True)
    for model in a.models:
        mk = dict(kw, data_model=model)
        if "tta" in a.what and model != "blobs":
            out[f"tta_{model}"] = ex.time_to_accuracy(0.10, blocks=a.blocks, **mk)
            print("TTA", model, json.dumps(out[f"tta_{model}"]), flush=True)
        if "poison" in a.what:
            for eps in (0.5, 2.0):
                key = f"poison30_eps{eps}_{model}"
                out[key] = ex.time_to_accuracy(0.10, blocks=a.blocks, poison_fraction=0.3, epsilon=eps, **mk)
                print("POISON", key, json.dumps(out[key]), flush=True)
        if "heatmap" in a.what:
            out[f"heatmap_{model}"] = ex.poison_heatmap(fractions=(0.4, 0.3, 0.1), collected=(40, 70), blocks=a.blocks, **mk)
            print("HEATMAP", model, json.dumps(out[f"heatmap_{model}"]), flush=True)
            if model == "lowrank":      # half the noise of the -ep default: where the defence starts to hold on this data
                out["heatmap_lowrank_eps4"] = ex.poison_heatmap(fractions=(0.4, 0.3, 0.1), collected=(40, 70), blocks=a.blocks,
                                                                epsilon=4.0, **mk)
                print("HEATMAP lowrank eps4", json.dumps(out["heatmap_lowrank_eps4"]), flush=True)
if "cifar" in a.what:
    out["cifar200"] = ex.time_to_accuracy(0.5, peers=200, dataset="cifar", blocks=50, shard_size=250, test_size=1000,
                                          data_sep=a.sep[0])
    print("CIFAR", json.dumps(out["cifar200"]), flush=True)
if "cifar_sgd" in a.what:
    from biscotti_b200.sim import ml_main

    for model, lr in (("blobs", 0.01), ("lowrank", 0.01)):
        out[f"cifar_sgd_{model}"] = ml_main.run_clients("cifar", honest=50, poisoners=0, rounds=300, optimizer="sgd", lr=lr, log_every=30,
                                                        device="cuda:0", data_model=model, shard_size=250, test_size=1000)
        print("CIFAR_SGD", model, json.dumps(out[f"cifar_sgd_{model}"]), flush=True)
out["wall_s"] = time.time() - t0
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/campaigns.json", "w"), indent=1)
