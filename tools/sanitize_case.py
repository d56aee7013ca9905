"""A small evaluation that reaches every kernel (many-pairs scan, fold-shared prediction, wide and narrow histogram cells,
sampling genes, both objectives), meant to be run under compute-sanitizer:
    compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gentun_b200.engine import Engine
from tests.synth import make_dataset

def genes(**kw):
    g = {'eta': 0.3, 'min_child_weight': 1, 'max_depth': 6, 'gamma': 0.0, 'max_delta_step': 0, 'subsample': 1.0,
         'colsample_bytree': 1.0, 'colsample_bylevel': 1.0, 'lambda': 1.0, 'alpha': 0.0, 'scale_pos_weight': 1.0}
    g.update(kw)
    return g

for f, cls in ((50, False), (70, True)):
    x, y = make_dataset(62000, f, seed=77 + f, classification=cls)
    eng = Engine(x, y)
    gs = [genes(max_depth=10, min_child_weight=0, eta=0.2), genes(max_depth=10, max_delta_step=2, alpha=0.3, subsample=0.6),
          genes(max_depth=4, colsample_bytree=0.5, colsample_bylevel=0.7), genes(max_depth=7, gamma=1.0, scale_pos_weight=2.0)]
    params = dict(nfold=5, num_boost_round=3)
    if cls:
        params.update(objective="binary:logistic", eval_metric="logloss")
    fit, rounds = eng.evaluate(gs, **params)
    print(f, cls, fit, rounds, flush=True)
    eng.close()
print("done")

# the shared-memory-resident regime (csrc/small.cu): team barriers, per-feature lists, early stopping, more
# individuals than teams is not needed here (the queue is exercised by tests/test_gpu_small.py)
x, y = make_dataset(3000, 9, seed=5)
eng = Engine(x, y)
fit, rounds = eng.evaluate([genes(max_depth=8, min_child_weight=0), genes(max_depth=5, subsample=0.7, colsample_bytree=0.6),
                            genes(max_depth=3, eta=1.0)], nfold=3, num_boost_round=12, early_stopping_rounds=3)
print("small", fit, rounds, eng.stats()["kernel_launches"], flush=True)
eng.close()
print("done small")
