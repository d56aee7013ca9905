#!/usr/bin/env python
"""Freezes outputs of the CPU oracle (oracle/gbt_cv_oracle.c) on small inputs into
tests/golden/oracle_wine.json, so that a change to the oracle is visible in review.

The reference's arithmetic lives in third-party xgboost, which is not installed here, so these are
NOT xgboost outputs: parity against real xgboost stays unpinned (see the oracle's header).  What is
pinned against an independent implementation is the fold permutation (numpy's legacy MT19937).

    python tests/golden/make_oracle_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402

GENES = {'eta': 0.3, 'min_child_weight': 1, 'max_depth': 6, 'gamma': 0.0, 'max_delta_step': 0, 'subsample': 1.0,
         'colsample_bytree': 1.0, 'colsample_bylevel': 1.0, 'lambda': 1.0, 'alpha': 0.0, 'scale_pos_weight': 1.0}


def main():
    data = pd.read_csv(os.path.join(ROOT, "tests", "data", "winequality-white.csv"), delimiter=";")
    y = data["quality"].values
    x = data.drop(["quality"], axis=1).values
    ctx = orc.Context(x, y, 3)
    out = {"fold_sha1": {}, "ncuts": {}, "cuts_sha1": {}, "bins_sha1": {}}
    for n, nfold in [(4898, 3), (1000000, 5), (10000000, 3), (4000000, 5)]:
        perm, fb = orc.make_folds(n, nfold)
        out["fold_sha1"]["%d/%d" % (n, nfold)] = {"sha1": hashlib.sha1(perm.tobytes()).hexdigest()[:12],
                                                   "head": perm[:5].tolist(), "sizes": np.diff(fb).tolist()}
    for k in range(3):
        cu, nc = ctx.cuts(k)
        out["ncuts"][str(k)] = nc.tolist()
        out["cuts_sha1"][str(k)] = hashlib.sha1(np.concatenate([cu[j, :nc[j]] for j in range(11)]).tobytes()).hexdigest()[:16]
        out["bins_sha1"][str(k)] = hashlib.sha1(ctx.train_bins(k).tobytes()).hexdigest()[:16]
    r = ctx.cv(GENES, orc.make_opts(nfold=3, num_boost_round=12), trace=True)
    out["trace_default_genes_12_rounds"] = [float(v) for v in r["trace"]]
    r = ctx.cv(GENES, orc.make_opts(nfold=3))
    out["default_genes_early_stop"] = {"fitness": r["fitness"], "rounds": r["rounds"], "best_iteration": r["best_iteration"]}
    g2 = dict(GENES, max_depth=4, max_delta_step=2, alpha=0.5, gamma=0.3, min_child_weight=3, eta=0.1, **{"lambda": 2.0})
    r = ctx.cv(g2, orc.make_opts(nfold=3, num_boost_round=10), trace=True)
    out["trace_regularised_10_rounds"] = [float(v) for v in r["trace"]]
    with open(os.path.join(HERE, "oracle_wine.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)
    print("wrote oracle_wine.json")


if __name__ == "__main__":
    main()
