import sys, os
sys.path.insert(0, '/root/repo')
import numpy as np
from gentun_b200.engine import Engine
from tests.synth import make_dataset, seeded_population_genes
x, y = make_dataset(1000000, 50, 1003)
genes = seeded_population_genes(256)
eng = Engine(x, y, collect_timing=True)
eng.prepare(5)
for s in range(2):
    eng.evaluate(genes[96+s*32:128+s*32], nfold=5, num_boost_round=10)
    st = eng.stats()
    print(os.environ.get('GT_HIST_DBG', '0'), 'eval %.0f ms hist %.1f ms  GB/s %.0f' % (st['eval_ms'], st['hist_ms'], st['hist_alg_bytes']/st['hist_ms']/1e6))
