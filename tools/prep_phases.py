"""Set-up phases of an engine (create + prepare) at the benchmark shapes, from the library's own trace
(GENTUN_B200_TRACE=1, stderr).  usage: python tools/prep_phases.py [c3|c4|c5 ...]"""
import os, sys, time
os.environ["GENTUN_B200_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gentun_b200.engine import Engine

SHAPES = {"c3": (1_000_000, 50, 5), "c4": (10_000_000, 100, 3), "c5": (4_000_000, 200, 5)}
for name in (sys.argv[1:] or ["c3"]):
    n, f, nfold = SHAPES[name]
    rng = np.random.default_rng(7)
    x = np.empty((f, n), dtype=np.float64)            # column-major float64, as a DataFrame holds it
    for j in range(f):
        x[j] = rng.standard_normal(n)
    y = (x[0] + 0.5 * x[1] > 0).astype(np.float64) if name == "c4" else x[0] * 2.0 + x[1]
    print("== %s: %d x %d float64, nfold %d" % (name, n, f, nfold), flush=True)
    for rep in range(2):
        t0 = time.perf_counter(); eng = Engine(x.T, y)
        t1 = time.perf_counter(); eng.prepare(nfold)
        t2 = time.perf_counter(); eng.close()
        t3 = time.perf_counter()
        print("rep %d: create %.0f ms, prepare %.0f ms, close %.0f ms" % (rep, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3), flush=True)
    del x, y
