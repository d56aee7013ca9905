import sys, os, time, random, contextlib, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pandas as pd
from gentun_b200.engine import Engine
from tests.synth import seeded_population_genes
d = pd.read_csv('tests/data/winequality-white.csv', delimiter=';')
y = d['quality'].values; x = d.drop(['quality'], axis=1).values
genes = seeded_population_genes(100)
eng = Engine(x, y)
eng.prepare(3)
rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 200
t = time.time()
fit, r = eng.evaluate(genes, nfold=3, num_boost_round=rounds, early_stopping_rounds=100)
dt = time.time() - t
st = eng.stats()
print("wine: 100 individuals, max %d rounds: %.2f s, %.1f us per launch, launches %d, rounds run max %d" % (rounds, dt, st['eval_ms']*1e3/st['kernel_launches'], st['kernel_launches'], r.max()))
