"""Recompute, in float64 numpy, the candidate the engine chose where its tree first leaves the oracle's."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gentun_b200.engine import Engine
from oracle import orc
from tests.synth import make_dataset
from tests.treediff import root_causes
d = json.load(open('tests/golden/drift_100_rounds.json'))['logistic_mcw']
R = 30
x, y = make_dataset(d['shape'][0], d['shape'][1], d['data_seed'], classification=True)
params = dict(d['params'], num_boost_round=R)
g = d['genes']
eng = Engine(x, y)
fit, rounds, extra = eng.evaluate([g], trace=True, tree_rounds=R, **params)
ctx = orc.Context(x, y, params['nfold'], nthread=orc.host_cores())
want = ctx.cv(g, orc.make_opts(nthread=orc.host_cores(), **params), trace=True, trees_rounds=R)
k = 0
n, c = root_causes(extra['trees'][0, k], want['trees'][k], g['gamma'])
r, nid, kind, margin = c[0]
print('first root cause: round', r, 'node', nid, kind, margin)
tr, te = eng.fold_rows(3, k)
bins = eng.bins(3, k)[:len(tr)]
yt = y[tr]
spw = np.float32(g['scale_pos_weight'])
w = np.where(yt == 1.0, spw, np.float32(1.0)).astype(np.float32)
sg = float(np.sum(((np.float32(0.5) - yt) * w).astype(np.float64))); sh = float(np.sum((np.float32(0.25) * w).astype(np.float64)))
stump = np.float32(-sg / max(sh, 1e-6))
bs = np.float32(1.0) / (np.exp(np.minimum(-stump, np.float32(88.7))) + np.float32(1.0) + np.float32(1e-16))
base = -np.log(np.float32(1.0) / bs - np.float32(1.0))
margin_v = np.full(len(tr), base, dtype=np.float32)
def leaf_of(tree, bins):
    nidv = np.zeros(len(bins), dtype=np.int64)
    for _ in range(11):
        f = tree['feat'][nidv]
        split = f >= 0
        if not split.any(): break
        b = bins[np.arange(len(bins)), np.where(split, f, 0)]
        nxt = 2 * nidv + 1 + (b > tree['bin'][nidv])
        nidv = np.where(split, nxt, nidv)
    return nidv
for rr in range(r):
    t = want['trees'][k, rr]
    margin_v = (margin_v + t['value'][leaf_of(t, bins)]).astype(np.float32)
p = (np.float32(1.0) / (np.exp(np.minimum(-margin_v, np.float32(88.7))) + np.float32(1.0) + np.float32(1e-16))).astype(np.float32)
gv = ((p - yt) * w).astype(np.float32); hv = (np.maximum(p * (np.float32(1.0) - p), np.float32(1e-16)) * w).astype(np.float32)
# rows of node nid under the ORACLE's tree r (identical above the root cause)
t = want['trees'][k, r]
te_ = extra['trees'][0, k, r]
path = []
q = nid
while q > 0:
    path.append(q); q = (q - 1) // 2
mask = np.ones(len(tr), dtype=bool)
q = 0
for child in reversed(path):
    b = bins[:, t['feat'][q]]
    mask &= (b <= t['bin'][q]) if child == 2 * q + 1 else (b > t['bin'][q])
    q = child
print('rows in node', mask.sum(), 'G', gv[mask].astype(np.float64).sum(), 'H', hv[mask].astype(np.float64).sum())
lam, mcw = float(g['lambda']), float(g['min_child_weight'])
def gain(G, H): return G * G / (H + lam) if H > 0 else 0.0
G, H = gv[mask].astype(np.float64).sum(), hv[mask].astype(np.float64).sum()
for name, (f, b) in (('engine', (te_['feat'][nid], te_['bin'][nid])), ('oracle', (t['feat'][nid], t['bin'][nid]))):
    left = mask & (bins[:, f] <= b)
    GL, HL = gv[left].astype(np.float64).sum(), hv[left].astype(np.float64).sum()
    print(name, 'split', f, b, 'left rows', left.sum(), 'GL', GL, 'HL', HL, 'GR', G - GL, 'HR', H - HL,
          'loss_chg', gain(GL, HL) + gain(G - GL, H - HL) - gain(G, H), 'valid', HL >= mcw and H - HL >= mcw)
sgq = 2.0 ** (19 - np.frexp(np.abs(gv).max())[1]); shq = 2.0 ** (17 - np.frexp(hv.max())[1])
gq = np.rint(gv.astype(np.float64) * sgq); hq = np.rint(hv.astype(np.float64) * shq)
f, b = te_['feat'][nid], te_['bin'][nid]
left = mask & (bins[:, f] <= b)
print('quantised: HL', hq[left].sum() / shq, 'HR', hq[mask & ~left].sum() / shq, 'GL', gq[left].sum() / sgq, 'GR', gq[mask & ~left].sum() / sgq, 'scales', sgq, shq)
print('engine dump node', te_[nid], 'children', te_[2 * nid + 1], te_[2 * nid + 2])
