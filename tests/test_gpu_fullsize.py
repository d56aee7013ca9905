"""BASELINE.json's shapes at (or near) full size, where the oracle is too slow to be the checker:
size-independent properties instead (SURVEY.md 8c / task rule 3)."""
import numpy as np
import pytest

from tests.synth import make_dataset, seeded_population_genes

pytestmark = pytest.mark.gpu


def _properties(n, f, nfold, classification, rounds=4, n_ind=3):
    from gentun_b200.engine import Engine
    x, y = make_dataset(n, f, seed=1000 + f, classification=classification)
    eng = Engine(x, y)
    # X1: folds partition the rows, sizes follow np.array_split
    np.random.seed(0)
    perm = np.random.permutation(n)
    sizes = [len(c) for c in np.array_split(perm, nfold)]
    seen = np.zeros(n, dtype=np.int32)
    for k in range(nfold):
        tr, te = eng.fold_rows(nfold, k)
        assert len(te) == sizes[k] and len(tr) == n - sizes[k]
        seen[te] += 1
        if k == 0:
            assert np.array_equal(te, perm[:sizes[0]])
    assert np.all(seen == 1)
    # X2/X3: strictly ascending cuts with the sentinel above the column maximum; bins are the upper bound
    for k in (0, nfold - 1):
        cu, nc = eng.cuts(nfold, k)
        tr, te = eng.fold_rows(nfold, k)
        bins = eng.bins(nfold, k)
        for j in range(0, f, max(1, f // 7)):
            c = cu[j, :nc[j]]
            assert 1 <= nc[j] <= 256 and np.all(np.diff(c) > 0)
            assert c[-1] > x[tr, j].max()
            rows = np.concatenate([tr, te])[::97]
            want = np.minimum(np.searchsorted(c, x[rows, j], side="right"), nc[j] - 1)
            assert np.array_equal(bins[::97, j], want.astype(np.uint8))
            # quantile property of the sketch: no bin of the train slice is grossly over-full
            counts = np.bincount(bins[:len(tr), j], minlength=nc[j])
            if nc[j] == 256:
                assert counts.max() <= 4 * len(tr) / 256
    # X4-X9: identical results on a second run, validation metric falling, deeper/shallower both fine
    genes = seeded_population_genes(64)
    genes = sorted(genes, key=lambda g: -g["max_depth"])[:1] + sorted(genes, key=lambda g: g["max_depth"])[:n_ind - 1]
    params = dict(nfold=nfold, num_boost_round=rounds)
    if classification:
        params.update(objective="binary:logistic", eval_metric="logloss")
    fit, r, extra = eng.evaluate(genes, trace=True, **params)
    fit2, _ = eng.evaluate(list(reversed(genes)), **params)       # other batch order, same answers
    assert np.array_equal(fit, fit2[::-1])
    for i in range(n_ind):
        t = extra["trace"][i, :rounds]
        assert np.all(np.isfinite(t)) and t[-1] < t[0]
    assert np.all(r == rounds)
    eng.close()


def test_c3_shape_parity_with_oracle():
    """The benchmark's own shape (1M x 50, nfold 5) against the oracle: cut points and bin indices bit-exact on every
    fold, per-round CV means within 1e-4 relative (north_star's fitness tolerance) for the deepest individual of the
    seeded population, one with max_delta_step != 0 and alpha != 0, and one with sampling genes, and tree identity
    graded node by node (tests/treediff.py)."""
    from gentun_b200.engine import Engine
    from oracle import orc
    from tests.treediff import grade, root_causes
    n, f, nfold, R = 1_000_000, 50, 5, 5
    x, y = make_dataset(n, f, seed=1003)
    eng = Engine(x, y)
    ctx = orc.Context(x, y, nfold, nthread=orc.host_cores())
    for k in range(nfold):
        cu, nc = eng.cuts(nfold, k)
        ocu, onc = ctx.cuts(k)
        assert np.array_equal(nc, onc)
        for j in range(f):
            assert np.array_equal(cu[j, :nc[j]].view(np.uint32), ocu[j, :nc[j]].view(np.uint32)), (k, j)
        if k in (0, nfold - 1):
            ntr = ctx.ntrain(k)
            assert np.array_equal(eng.bins(nfold, k)[:ntr], ctx.train_bins(k))     # oracle's own C binning of the train slice
    pop = seeded_population_genes(256)
    deep = next(g for g in pop if g["max_depth"] == 10 and g["subsample"] == 1.0 and g["colsample_bytree"] == 1.0 and g["colsample_bylevel"] == 1.0)
    reg = next(g for g in pop if g["max_delta_step"] != 0 and g["alpha"] != 0 and g["subsample"] == 1.0
               and g["colsample_bytree"] == 1.0 and g["colsample_bylevel"] == 1.0 and g is not deep)
    samp = next(g for g in pop if g["subsample"] < 1.0 and g["colsample_bytree"] < 1.0)
    gs = [deep, reg, samp]
    fit, rounds, extra = eng.evaluate(gs, trace=True, tree_rounds=R, nfold=nfold, num_boost_round=R)
    compared, causes = 0, []
    for i, g in enumerate(gs):
        want = ctx.cv(g, orc.make_opts(nfold=nfold, num_boost_round=R, nthread=orc.host_cores()), trace=True, trees_rounds=R)
        np.testing.assert_allclose(extra["trace"][i, :R], want["trace"], rtol=1e-4)       # north_star: 1e-4 relative
        assert abs(fit[i] - want["fitness"]) <= 1e-4 * abs(want["fitness"])
        for k in range(nfold):
            n_cmp, c = root_causes(extra["trees"][i, k], want["trees"][k], g["gamma"])
            compared += n_cmp
            causes += c
    ties, flips, wrong = grade(causes)
    print("tree identity at 1M x 50: %d nodes compared, %d root-cause mismatches (%d fp32 ties, %d quantisation flips)"
          % (compared, len(causes), ties, flips))
    assert not wrong, wrong
    assert len(causes) <= 0.01 * compared, (len(causes), compared)
    eng.close()
    ctx.close()


def test_hundred_round_drift_against_frozen_oracle():
    """Full-length fits (100 rounds) against oracle traces frozen by tests/golden/make_drift_golden.py: the benchmark
    shape under squared error, and a logistic fit with min_child_weight near the deep nodes' hessian sums.  Every round's
    CV mean stays within north_star's 1e-4 relative; the recorded drift is quoted in DESIGN.md."""
    import json
    import os
    from gentun_b200.engine import Engine
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "drift_100_rounds.json")))
    for name, c in gold.items():
        x, y = make_dataset(c["shape"][0], c["shape"][1], c["data_seed"], classification=c["classification"])
        eng = Engine(x, y)
        fit, rounds, extra = eng.evaluate([c["genes"]], trace=True, **c["params"])
        eng.close()
        want = np.asarray(c["oracle_trace"])
        got = extra["trace"][0, :len(want)]
        np.testing.assert_allclose(got, want, rtol=1e-4, err_msg=name)      # north_star: 1e-4 relative, every round
        assert rounds[0] == len(want)


@pytest.mark.parametrize("n,f,nfold,cls", [(2_000_000, 100, 3, True), (1_000_000, 200, 5, False)])
def test_c4_c5_feature_widths_parity_with_oracle(n, f, nfold, cls):
    """The other two streaming shapes against the oracle at a size it finishes in seconds: 100 features (four groups, pair
    CTAs, logistic / logloss as in C4) and 200 features (seven groups, single-group CTAs, as in C5); every round's CV
    mean within 1e-4 for a deep and a regularised individual."""
    from gentun_b200.engine import Engine
    from oracle import orc
    x, y = make_dataset(n, f, seed=1000 + f, classification=cls)
    eng = Engine(x, y)
    ctx = orc.Context(x, y, nfold, nthread=orc.host_cores())
    pop = seeded_population_genes(64)
    plain = [g for g in pop if g["subsample"] == 1.0 and g["colsample_bytree"] == 1.0 and g["colsample_bylevel"] == 1.0]
    gs = [max(plain, key=lambda g: g["max_depth"]), next(g for g in plain if g["max_delta_step"] != 0 and g["max_depth"] <= 6)]
    params = dict(nfold=nfold, num_boost_round=3)
    if cls:
        params.update(objective="binary:logistic", eval_metric="logloss")
    fit, rounds, extra = eng.evaluate(gs, trace=True, **params)
    for i, g in enumerate(gs):
        want = ctx.cv(g, orc.make_opts(nthread=orc.host_cores(), **params), trace=True)
        np.testing.assert_allclose(extra["trace"][i, :3], want["trace"], rtol=1e-4)      # north_star: 1e-4 relative
    eng.close()
    ctx.close()


def test_c3_shape_1m_by_50_regression():
    _properties(1_000_000, 50, 5, False)


def test_c4_shape_logistic_100_features():
    _properties(10_000_000, 100, 3, True, rounds=3, n_ind=2)       # C4 at its full 10M rows (4 GB of host data)


def test_c5_shape_200_features():
    _properties(1_000_000, 200, 5, False)      # C5 is 4M rows; 7 feature groups (odd: single-group CTAs)
