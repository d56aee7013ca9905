import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gentun_b200.engine import Engine
from tests.synth import make_dataset, seeded_population_genes
x, y = make_dataset(1000000, 50, 1003)
genes = seeded_population_genes(256)
for rep in range(3):
    t0 = time.perf_counter(); eng = Engine(x, y)
    t1 = time.perf_counter(); eng.prepare(5)
    t2 = time.perf_counter(); eng.evaluate(genes[rep*32:(rep+1)*32], nfold=5, num_boost_round=100)
    t3 = time.perf_counter(); dev = eng.stats()['eval_ms']; eng.close()
    t4 = time.perf_counter()
    print("create %.0f ms  prepare %.0f ms  eval %.0f ms (device %.0f)  close %.0f ms" % ((t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, dev, (t4-t3)*1e3))
