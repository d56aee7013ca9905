import sys, time, os
sys.path.insert(0, '/root/repo')
import numpy as np
from gentun_b200.engine import Engine
from tests.synth import make_dataset, seeded_population_genes
x, y = make_dataset(1000000, 50, 1003)
genes = seeded_population_genes(256)
for timing in (False, True):
    eng = Engine(x, y, collect_timing=timing)
    eng.prepare(5)
    for s in range(4):
        t = time.perf_counter()
        eng.evaluate(genes[s*8:(s+1)*8], nfold=5, num_boost_round=100)
        w = (time.perf_counter() - t) * 1e3
        print('timing', timing, 'step', s, 'wall %.0f ms device %.0f ms launches %d' % (w, eng.stats()['eval_ms'], eng.stats()['kernel_launches']), flush=True)
    eng.close()
