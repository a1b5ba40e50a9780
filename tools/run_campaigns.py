#!/usr/bin/env python
"""Full-length campaigns on one GPU (results -> gpurun_out/campaigns.json, copied to profiles/ by hand):

  * time to accuracy, MNIST-shaped softmax, 100 peers, 102 blocks, reference defaults (parseLogs.py:197-210)
  * the same with 30 % poisoners and epsilon = 0.5 (BASELINE.json config 3)
  * poisoning heat-map rows with -rs=true (plotHeatMap.py:46-88)
  * CIFAR-shaped CNN, 200 peers: does the test error fall over 50 blocks

The synthetic classes overlap (`data_sep` = 3.6, the package default) so that the error needs tens of blocks to pass 0.10."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.eval import experiments as ex  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--sep", type=float, nargs="*", default=[3.6])
ap.add_argument("--what", nargs="*", default=["tta", "poison", "heatmap", "cifar"])
ap.add_argument("--blocks", type=int, default=102)
a = ap.parse_args()
out = {}
t0 = time.time()
for sep in a.sep:
    kw = dict(data_sep=sep)
    if "tta" in a.what:
        out[f"tta_sep{sep}"] = ex.time_to_accuracy(0.10, blocks=a.blocks, **kw)
        print("TTA", sep, json.dumps(out[f"tta_sep{sep}"]), flush=True)
    if "poison" in a.what:
        out[f"poison30_eps0.5_sep{sep}"] = ex.time_to_accuracy(0.10, blocks=a.blocks, poison_fraction=0.3, epsilon=0.5, **kw)
        print("POISON", sep, json.dumps(out[f"poison30_eps0.5_sep{sep}"]), flush=True)
    if "heatmap" in a.what:
        out[f"heatmap_sep{sep}"] = ex.poison_heatmap(fractions=(0.4, 0.3, 0.1), collected=(40, 70), blocks=a.blocks, **kw)
        print("HEATMAP", sep, json.dumps(out[f"heatmap_sep{sep}"]), flush=True)
if "cifar" in a.what:
    out["cifar200"] = ex.time_to_accuracy(0.5, peers=200, dataset="cifar", blocks=50, shard_size=250, test_size=1000,
                                          data_sep=a.sep[0])
    print("CIFAR", json.dumps(out["cifar200"]), flush=True)
out["wall_s"] = time.time() - t0
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/campaigns.json", "w"), indent=1)
