"""Per-round time of the shared-memory-resident kernel on winequality, by tree depth and by how many jobs share the GPU."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pandas as pd
from gentun_b200.engine import Engine
from tests.test_gpu_parity import genes
df = pd.read_csv('tests/data/winequality-white.csv', delimiter=';')
x = np.ascontiguousarray(df.drop(['quality'], axis=1).values, dtype=np.float32); y = np.ascontiguousarray(df['quality'].values, dtype=np.float32)
eng = Engine(x, y); eng.prepare(3)
R = 300
for copies in (1, 100):
    for d in (3, 6, 10):
        gs = [genes(max_depth=d, eta=0.01, min_child_weight=0 if d == 10 else 1) for _ in range(copies)]
        for i, g in enumerate(gs): g['gamma'] = 1e-4 * i     # distinct gene sets
        eng.evaluate(gs, nfold=3, num_boost_round=R, early_stopping_rounds=0)
        print('copies %3d depth %2d: %.1f us per round (queue depth %d)' % (copies, d, eng.stats()['eval_ms'] * 1e3 / R, eng.stats()['waves']))
