import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gentun_b200.engine import Engine
from tests.synth import make_dataset, seeded_population_genes

n, f, nfold, cls = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
rounds = 4
x, y = make_dataset(n, f, seed=1004, classification=bool(cls))
eng = Engine(x, y)
genes = seeded_population_genes(64)
genes = sorted(genes, key=lambda g: -g["max_depth"])[:1]
params = dict(nfold=nfold, num_boost_round=rounds)
if cls:
    params.update(objective="binary:logistic", eval_metric="logloss")
runs = []
for rep in range(3):
    fit, r, extra = eng.evaluate(genes, trace=True, tree_rounds=rounds, **params)
    runs.append((fit.copy(), extra["trace"].copy(), extra["trees"].copy()))
    print("run", rep, fit, extra["trace"][0, :rounds])
a = runs[0]
for rep in (1, 2):
    b = runs[rep]
    if np.array_equal(a[0], b[0]):
        print("run", rep, "identical")
        continue
    for k in range(nfold):
        for r in range(rounds):
            ta, tb = a[2][0, k, r], b[2][0, k, r]
            diff = np.nonzero((ta["feat"] != tb["feat"]) | (ta["bin"] != tb["bin"]) | (ta["value"] != tb["value"]))[0]
            if len(diff):
                nid = diff[0]
                print("first diff: fold", k, "round", r, "node", nid, "depth", int(np.log2(nid + 1)), ta[nid], tb[nid], "n diff", len(diff))
                break
        else:
            continue
        break
