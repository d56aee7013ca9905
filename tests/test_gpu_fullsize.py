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


def test_c3_shape_1m_by_50_regression():
    _properties(1_000_000, 50, 5, False)


def test_c4_shape_logistic_100_features():
    _properties(2_000_000, 100, 3, True)       # C4's 10M rows need 4 GB of host data; same kernels, same groups


def test_c5_shape_200_features():
    _properties(1_000_000, 200, 5, False)      # C5 is 4M rows; 7 feature groups (odd: single-group CTAs)
