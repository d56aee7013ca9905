"""Freezes full-length per-round CV traces of the oracle (and of the engine beside them) so that the drift of the
engine's 2^-19 gradient quantisation over a whole 100-round fit is a recorded number, and so that the GPU suite can
check 100 rounds against the oracle without paying the oracle's minutes every time.

Run on a GPU box from the repo root:  python tests/golden/make_drift_golden.py [out.json]
(two cases: the C3 benchmark shape, squared error; a 200 000 x 30 binary:logistic fit whose min_child_weight sits
near the hessian sums of the deep nodes).  tests/test_gpu_fullsize.py::test_hundred_round_drift_against_frozen_oracle
reads the result from tests/golden/drift_100_rounds.json.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

from oracle import orc  # noqa: E402
from tests.synth import make_dataset, seeded_population_genes  # noqa: E402

CASES = {
    "c3_regression": dict(n=1_000_000, f=50, seed=1003, cls=False, nfold=5, rounds=100,
                          pick=lambda pop: next(g for g in pop if g["max_depth"] == 7 and g["subsample"] == 1.0 and
                                                g["colsample_bytree"] == 1.0 and g["colsample_bylevel"] == 1.0)),
    "logistic_mcw": dict(n=200_000, f=30, seed=1004, cls=True, nfold=3, rounds=100,
                         pick=lambda pop: {'eta': 0.3, 'min_child_weight': 8, 'max_depth': 8, 'gamma': 0.0, 'max_delta_step': 0,
                                           'subsample': 1.0, 'colsample_bytree': 1.0, 'colsample_bylevel': 1.0, 'lambda': 1.0,
                                           'alpha': 0.0, 'scale_pos_weight': 2.0}),
}


def main():
    from gentun_b200.engine import Engine
    out = {}
    pop = seeded_population_genes(256)
    for name, c in CASES.items():
        x, y = make_dataset(c["n"], c["f"], c["seed"], classification=c["cls"])
        g = c["pick"](pop)
        params = dict(nfold=c["nfold"], num_boost_round=c["rounds"], early_stopping_rounds=0)
        if c["cls"]:
            params.update(objective="binary:logistic", eval_metric="logloss")
        eng = Engine(x, y)
        fit, rounds, extra = eng.evaluate([g], trace=True, **params)
        eng.close()
        ctx = orc.Context(x, y, c["nfold"], nthread=orc.host_cores())
        want = ctx.cv(g, orc.make_opts(nthread=orc.host_cores(), **params), trace=True)
        ctx.close()
        gpu = extra["trace"][0, :c["rounds"]]
        rel = np.abs(gpu - want["trace"]) / np.abs(want["trace"])
        out[name] = {"genes": g, "params": params, "shape": [c["n"], c["f"]], "data_seed": c["seed"], "classification": c["cls"],
                     "oracle_trace": want["trace"].tolist(), "engine_trace_when_frozen": gpu.tolist(),
                     "max_rel_err": float(rel.max()), "rel_err_last_round": float(rel[-1])}
        print(name, "max rel err over %d rounds: %.3e (last round %.3e)" % (c["rounds"], rel.max(), rel[-1]))
    path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "drift_100_rounds.json")
    json.dump(out, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
