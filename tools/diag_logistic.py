"""Where does the engine's logistic fit leave the oracle's?  (diagnostic for tests/golden/drift_100_rounds.json)"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gentun_b200.engine import Engine
from oracle import orc
from tests.synth import make_dataset
from tests.treediff import root_causes
d = json.load(open('tests/golden/drift_100_rounds.json'))['logistic_mcw']
R = int(os.environ.get('R', '40'))
x, y = make_dataset(d['shape'][0], d['shape'][1], d['data_seed'], classification=True)
params = dict(d['params'], num_boost_round=R)
eng = Engine(x, y)
fit, rounds, extra = eng.evaluate([d['genes']], trace=True, tree_rounds=R, **params)
ctx = orc.Context(x, y, params['nfold'], nthread=orc.host_cores())
want = ctx.cv(d['genes'], orc.make_opts(nthread=orc.host_cores(), **params), trace=True, trees_rounds=R)
a = want['trace']; b = extra['trace'][0, :R]
rel = np.abs(a - b) / np.abs(a)
print(os.environ.get('GENTUN_B200_LIB', 'default'), 'first rounds with rel > 1e-6:', np.nonzero(rel > 1e-6)[0][:5], 'max rel', rel.max())
for k in range(params['nfold']):
    n, c = root_causes(extra['trees'][0, k], want['trees'][k], d['genes']['gamma'])
    print('fold', k, 'compared', n, 'root causes', c[:4])
    for (r, nid, kind, margin) in c[:2]:
        m, o = extra['trees'][0, k, r], want['trees'][k, r]
        for q in (nid, 2 * nid + 1, 2 * nid + 2, (nid - 1) // 2):
            if 0 <= q < 2047:
                print('   round', r, 'node', q, 'engine', m[q], 'oracle', o[q])
