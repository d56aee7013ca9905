import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pandas as pd
from gentun_b200.engine import Engine
from tests.test_gpu_parity import genes
df = pd.read_csv('tests/data/winequality-white.csv', delimiter=';')
x = np.ascontiguousarray(df.drop(['quality'], axis=1).values, dtype=np.float32); y = np.ascontiguousarray(df['quality'].values, dtype=np.float32)
eng = Engine(x, y); eng.prepare(3)
gs = [genes(max_depth=int(os.environ.get('DEPTH', '6')), eta=0.01, min_child_weight=int(os.environ.get('MCW', '1')), gamma=1e-4 * i) for i in range(int(os.environ.get('COPIES', '1')))]
eng.evaluate(gs, nfold=3, num_boost_round=int(os.environ.get('R', '30')), early_stopping_rounds=0)
print(eng.stats()['eval_ms'])
