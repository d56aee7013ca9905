"""Collects device times of random batches of the seeded C3 population (tools only; feeds sharding.fit_cost).
Writes gpurun_out/cost_samples.json: [{"idx": [...], "ms": device time}]."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gentun_b200.engine import Engine
from tests.synth import make_dataset, seeded_population_genes

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 25
batches = int(sys.argv[2]) if len(sys.argv) > 2 else 72
width = 16
x, y = make_dataset(1000000, 50, 1003)
genes = seeded_population_genes(256)
eng = Engine(x, y)
eng.prepare(5)
rng = np.random.default_rng(7)
out = []
eng.evaluate(genes[:width], nfold=5, num_boost_round=3)
for b in range(batches):
    if b < 16:      # homogeneous batches first: sorted by depth, then by subsample
        key = (lambda i: (genes[i]["max_depth"], genes[i]["subsample"])) if b < 8 else (lambda i: (genes[i]["subsample"], genes[i]["max_depth"]))
        order = sorted(range(256), key=key)
        idx = order[(b % 8) * 32:(b % 8) * 32 + 32:2]
    else:
        idx = rng.choice(256, size=width, replace=False).tolist()
    eng.evaluate([genes[i] for i in idx], nfold=5, num_boost_round=rounds)
    out.append({"idx": [int(i) for i in idx], "ms": eng.stats()["eval_ms"]})
    print(b, out[-1]["ms"], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"rounds": rounds, "samples": out, "genes": genes}, open("gpurun_out/cost_samples.json", "w"))
