"""Times one cost-balanced shard of the C3 population for a few rounds (kernel experiments; GENTUN_B200_LIB picks the build)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gentun_b200.engine import Engine
from gentun_b200.sharding import fit_cost, shard_plan
from tests.synth import make_dataset, seeded_population_genes
rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 10
x, y = make_dataset(1000000, 50, 1003)
genes = seeded_population_genes(256)
shard = shard_plan([fit_cost(g) for g in genes], 8)[0]
eng = Engine(x, y, collect_timing=True)
eng.prepare(5)
for s in range(2):
    fit, _ = eng.evaluate([genes[i] for i in shard], nfold=5, num_boost_round=rounds)
    st = eng.stats()
    print(os.path.basename(os.environ.get('GENTUN_B200_LIB', 'default')), 'eval %.1f ms  hist %.1f ms  fit0 %.9f' % (st['eval_ms'], st['hist_ms'], fit[0]), flush=True)
eng.close()
