"""Full-size robustness check of the non-bench configs (C4: 10M x 100 logistic, C5: 4M x 200) through
size-independent properties: ascending cuts, bins below ncuts and monotone in the value, identical
results on a second run, validation metric falling over the first rounds."""
import sys
import time
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gentun_b200.engine import Engine
from tests.synth import make_dataset, seeded_population_genes


def check(name, n, f, nfold, cls, n_ind, rounds):
    t = time.time()
    x, y = make_dataset(n, f, seed=1004, classification=cls)
    print(name, "data %.1fs" % (time.time() - t), flush=True)
    t = time.time()
    eng = Engine(x, y)
    eng.prepare(nfold)
    print(name, "create+prepare %.1fs (device %.0f ms)" % (time.time() - t, eng.stats()["prepare_ms"]), flush=True)
    cu, nc = eng.cuts(nfold, 0)
    for j in range(f):
        assert nc[j] >= 1 and np.all(np.diff(cu[j, :nc[j]]) > 0), j
    tr, te = eng.fold_rows(nfold, 0)
    assert len(set(tr[:1000]) & set(te[:1000])) == 0 and len(tr) + len(te) == n
    genes = seeded_population_genes(64)
    genes = sorted(genes, key=lambda g: -g["max_depth"])[:1] + sorted(genes, key=lambda g: g["max_depth"])[:n_ind - 1]
    params = dict(nfold=nfold, num_boost_round=rounds)
    if cls:
        params.update(objective="binary:logistic", eval_metric="logloss")
    t = time.time()
    fit, r, extra = eng.evaluate(genes, trace=True, **params)
    dt = time.time() - t
    fit2, _ = eng.evaluate(genes, **params)
    assert np.array_equal(fit, fit2), (fit, fit2)
    for i in range(n_ind):
        tr_ = extra["trace"][i, :rounds]
        assert np.all(np.isfinite(tr_)) and tr_[-1] <= tr_[0] + 1e-12, tr_
    st = eng.stats()
    print(name, "eval %.2fs depths %s fitness %s waves %d hist GB/s %.0f" % (
        dt, [g["max_depth"] for g in genes], np.round(fit, 5).tolist(), st["waves"],
        st["hist_alg_bytes"] / max(st["hist_ms"], 1e-9) / 1e6), flush=True)
    eng.close()


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("c4", "all"):
        check("C4", 10_000_000, 100, 3, True, 3, 4)
    if which in ("c5", "all"):
        check("C5", 4_000_000, 200, 5, False, 3, 4)
