"""Kernel experiments on the C3 shape: one shard of the seeded population for a few rounds with the library named by
GENTUN_B200_LIB (default: the in-tree build).  Prints eval / histogram times and a hash of the fitness vector --
integer histograms make every correct variant bit-equal, so the hash doubles as a regression check.

    python tools/exp.py [--rounds 10] [--reps 2] [--workload c3|c4|c5] [--inds 32]
"""
import argparse
import hashlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from gentun_b200.engine import Engine, LIB_PATH  # noqa: E402
from gentun_b200.sharding import fit_cost, shard_plan  # noqa: E402
from tests.synth import make_dataset, seeded_population_genes  # noqa: E402

SHAPES = {"c3": (1_000_000, 50, 5, False, 1003), "c3s": (200_000, 50, 5, False, 1003),
          "c4": (10_000_000, 100, 3, True, 1004), "c4s": (2_000_000, 100, 3, True, 1004),
          "c5": (4_000_000, 200, 5, False, 1005), "c5s": (1_000_000, 200, 5, False, 1005)}

ap = argparse.ArgumentParser()
ap.add_argument("--rounds", type=int, default=10)
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--workload", default="c3")
ap.add_argument("--inds", type=int, default=32)
ap.add_argument("--tag", default="")
a = ap.parse_args()

n, f, nfold, cls, seed = SHAPES[a.workload]
x, y = make_dataset(n, f, seed, classification=cls)
genes = seeded_population_genes(256)
shard = shard_plan([fit_cost(g) for g in genes], 256 // a.inds)[0]
gs = [genes[i] for i in shard]
params = dict(nfold=nfold, num_boost_round=a.rounds)
if cls:
    params.update(objective="binary:logistic", eval_metric="logloss")
eng = Engine(x, y, collect_timing=True)
eng.prepare(nfold)
out = {"lib": os.path.basename(LIB_PATH), "tag": a.tag, "workload": a.workload, "rounds": a.rounds}
for r in range(a.reps):
    fit, _ = eng.evaluate(gs, **params)
    st = eng.stats()
    out.update(eval_ms=round(st["eval_ms"], 1), hist_ms=round(st["hist_ms"], 1),
               hist_gbs=round(st["hist_alg_bytes"] / st["hist_ms"] / 1e6, 0), frac=round(st["hist_alg_bytes"] / st["hist_ms"] / 1e6 / 6560.6, 4),
               launches=st["kernel_launches"], fit_hash=hashlib.sha1(np.ascontiguousarray(fit).tobytes()).hexdigest()[:12])
print(json.dumps(out))
