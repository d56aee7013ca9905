"""The CPU oracle against what can be pinned without xgboost: numpy's fold stream, the frozen fixtures,
hand-checkable small cases and the early-stopping rule re-derived from its own per-round trace."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import orc

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "oracle_wine.json")))


def genes(**kw):
    g = {'eta': 0.3, 'min_child_weight': 1, 'max_depth': 6, 'gamma': 0.0, 'max_delta_step': 0, 'subsample': 1.0,
         'colsample_bytree': 1.0, 'colsample_bylevel': 1.0, 'lambda': 1.0, 'alpha': 0.0, 'scale_pos_weight': 1.0}
    g.update(kw)
    return g


@pytest.fixture(scope="module")
def wine_ctx(wine):
    x, y = wine
    ctx = orc.Context(x.values, y.values, 3)
    yield ctx
    ctx.close()


@pytest.mark.parametrize("n,nfold", [(4898, 3), (53, 5), (100003, 4), (2, 2)])
def test_folds_match_numpy_legacy_stream(n, nfold):
    perm, fb = orc.make_folds(n, nfold)
    np.random.seed(0)
    ref = np.random.permutation(n)
    assert np.array_equal(perm, ref)
    assert np.diff(fb).tolist() == [len(c) for c in np.array_split(ref, nfold)]


def test_fold_hashes_of_the_benchmark_shapes():
    # SURVEY.md A.1 recorded these with numpy 2.3.5
    want = {"4898/3": "0282d82433a1", "1000000/5": "dfa9fc2c6f01"}
    for key, sha in want.items():
        n, nfold = (int(v) for v in key.split("/"))
        perm, _ = orc.make_folds(n, nfold)
        assert hashlib.sha1(perm.tobytes()).hexdigest()[:12] == sha == GOLD["fold_sha1"][key]["sha1"]


def test_train_slice_order_is_concatenation_of_other_chunks(wine_ctx):
    np.random.seed(0)
    chunks = np.array_split(np.random.permutation(4898), 3)
    for k in range(3):
        assert np.array_equal(wine_ctx.test_rows(k), chunks[k])
        assert np.array_equal(wine_ctx.train_rows(k), np.concatenate([chunks[i] for i in range(3) if i != k]))


def test_cuts_of_small_columns_are_the_distinct_values(wine_ctx, wine):
    """Fewer rows than the sketch queue and <= 256 distinct values: the summary is exact, so the cuts are
    every distinct value but the smallest, then the sentinel max + (|max| + 1e-5)."""
    x = wine[0].values.astype(np.float32)
    for k in range(3):
        cu, nc = wine_ctx.cuts(k)
        rows = wine_ctx.train_rows(k)
        for j in range(11):
            vals = np.unique(x[rows, j])
            c = cu[j, :nc[j]]
            assert np.all(np.diff(c) > 0)
            sentinel = np.float32(vals[-1] + (np.abs(vals[-1]) + np.float32(1e-5)))
            assert c[-1] == sentinel
            if len(vals) <= 256:
                assert np.array_equal(c[:-1], vals[1:])
            else:
                assert nc[j] <= 256 and set(c[:-1]).issubset(set(vals))


def test_sketch_cuts_keep_the_quantile_guarantee():
    """A column with far more rows than the sketch queue (several flushes, prune / combine through the summary levels):
    WQuantileSketch promises ranks within eps * n of the requested quantiles with eps = 1 / (8 * max_bin), so the cut
    points must be data values and no bin may hold much more than n / 256 rows -- a property the restatement cannot meet
    by accident if its prune or combine were wrong.  Continuous, heavily tied and skewed columns."""
    rng = np.random.default_rng(11)
    n = 300_000
    x = np.empty((n, 3), dtype=np.float32)
    x[:, 0] = rng.standard_normal(n)
    x[:, 1] = np.round(rng.standard_normal(n) * 40.0)            # ~300 distinct values, heavy ties
    x[:, 2] = rng.lognormal(0.0, 2.0, n)                         # long right tail
    y = x[:, 0].copy()
    ctx = orc.Context(x, y, 2)
    try:
        for k in range(2):
            cu, nc = ctx.cuts(k)
            rows = ctx.train_rows(k)
            for j in range(3):
                col = x[rows, j]
                c = cu[j, :nc[j]]
                assert 2 <= nc[j] <= 256 and np.all(np.diff(c) > 0)
                assert set(c[:-1]).issubset(set(np.unique(col))) and c[-1] > col.max()
                counts = np.bincount(np.minimum(np.searchsorted(c, col, side="right"), nc[j] - 1), minlength=nc[j])
                if j == 1:
                    # ties: a bin cannot be split below one distinct value
                    biggest_tie = np.unique(col, return_counts=True)[1].max()
                    assert counts.max() <= biggest_tie + 2 * len(col) / 256
                else:
                    assert counts.max() <= 2.0 * len(col) / 256, (j, counts.max(), len(col) / 256)
    finally:
        ctx.close()


def test_bins_are_upper_bound_of_cuts(wine_ctx, wine):
    x = wine[0].values.astype(np.float32)
    cu, nc = wine_ctx.cuts(1)
    b = wine_ctx.train_bins(1)
    rows = wine_ctx.train_rows(1)
    for j in range(11):
        want = np.minimum(np.searchsorted(cu[j, :nc[j]], x[rows, j], side="right"), nc[j] - 1)
        assert np.array_equal(b[:, j], want.astype(np.uint8))
    allb = wine_ctx.all_bins(1)
    assert np.array_equal(allb[:len(rows)], b)


def test_frozen_fixtures(wine_ctx):
    for k in range(3):
        cu, nc = wine_ctx.cuts(k)
        assert nc.tolist() == GOLD["ncuts"][str(k)]
        h = hashlib.sha1(np.concatenate([cu[j, :nc[j]] for j in range(11)]).tobytes()).hexdigest()[:16]
        assert h == GOLD["cuts_sha1"][str(k)]
        assert hashlib.sha1(wine_ctx.train_bins(k).tobytes()).hexdigest()[:16] == GOLD["bins_sha1"][str(k)]
    r = wine_ctx.cv(genes(), orc.make_opts(nfold=3, num_boost_round=12), trace=True)
    np.testing.assert_allclose(r["trace"], GOLD["trace_default_genes_12_rounds"], rtol=1e-12)
    g2 = genes(max_depth=4, max_delta_step=2, alpha=0.5, gamma=0.3, min_child_weight=3, eta=0.1, **{"lambda": 2.0})
    r = wine_ctx.cv(g2, orc.make_opts(nfold=3, num_boost_round=10), trace=True)
    np.testing.assert_allclose(r["trace"], GOLD["trace_regularised_10_rounds"], rtol=1e-12)


def test_early_stopping_rule_rederived_from_trace(wine_ctx):
    """callback.EarlyStopping: strict improvement resets the counter; stop once `rounds` rounds passed
    without one; cv() truncates to best_iteration + 1 and gentun returns the last row."""
    esr = 7
    full = wine_ctx.cv(genes(eta=0.8, max_depth=8), orc.make_opts(nfold=3, num_boost_round=120, early_stopping_rounds=0), trace=True)
    got = wine_ctx.cv(genes(eta=0.8, max_depth=8), orc.make_opts(nfold=3, num_boost_round=120, early_stopping_rounds=esr), trace=True)
    best, best_it, stop_at = None, 0, None
    for it, m in enumerate(full["trace"]):
        if it == 0 or m < best:
            best, best_it = m, it
        elif it - best_it >= esr:
            stop_at = it
            break
    assert stop_at is not None and got["stopped_early"]
    assert got["rounds"] == stop_at + 1 and got["best_iteration"] == best_it
    assert got["fitness"] == best
    np.testing.assert_array_equal(got["trace"], full["trace"][:stop_at + 1])
    # without stopping, the fitness is the final round's mean, not the best
    assert full["fitness"] == full["trace"][-1] and not full["stopped_early"]
    r = wine_ctx.cv(genes(), orc.make_opts(nfold=3))
    assert (r["rounds"], r["best_iteration"]) == (GOLD["default_genes_early_stop"]["rounds"], GOLD["default_genes_early_stop"]["best_iteration"])
    assert r["fitness"] == GOLD["default_genes_early_stop"]["fitness"]


def _stump_cv_by_hand(x, y, nfold, lam, eta, mcw):
    """Independent numpy restatement for one round of depth-1 trees on data with few distinct values."""
    np.random.seed(0)
    chunks = np.array_split(np.random.permutation(len(y)), nfold)
    scores = []
    for k in range(nfold):
        te = chunks[k]
        tr = np.concatenate([chunks[i] for i in range(nfold) if i != k])
        base = np.float32(np.mean(y[tr].astype(np.float64)))
        g = (base - y[tr]).astype(np.float32).astype(np.float64)
        G, H = g.sum(), float(len(tr))
        best = (0.0, None)
        for j in range(x.shape[1]):
            for thr in np.unique(x[tr, j])[1:]:
                left = x[tr, j] < thr
                GL, HL = g[left].sum(), float(left.sum())
                GR, HR = G - GL, H - HL
                if HL < mcw or HR < mcw:
                    continue
                gain = np.float32(np.float32(GL * GL / (HL + lam)) + np.float32(GR * GR / (HR + lam))) - np.float32(G * G / (H + lam))
                if gain > best[0]:
                    best = (gain, (j, thr, GL, HL, GR, HR))
        pred = np.full(len(te), base, dtype=np.float32)
        if best[1] is not None and best[0] > 1e-6:
            j, thr, GL, HL, GR, HR = best[1]
            wl = np.float32(eta) * np.float32(-GL / (HL + lam))
            wr = np.float32(eta) * np.float32(-GR / (HR + lam))
            pred = pred + np.where(x[te, j] < thr, wl, wr).astype(np.float32)
        else:
            pred = pred + np.float32(eta) * np.float32(-G / (H + lam))
        d = (y[te] - pred).astype(np.float32)
        scores.append(np.sqrt(np.sum((d * d).astype(np.float64)) / len(te)))
    return float(np.mean(scores))


def test_one_round_of_stumps_matches_hand_computation():
    rng = np.random.default_rng(3)
    x = rng.integers(0, 6, size=(400, 3)).astype(np.float32)
    y = (x[:, 0] * 1.5 - (x[:, 1] > 2) * 2.0 + rng.standard_normal(400) * 0.3).astype(np.float32)
    ctx = orc.Context(x, y, 4)
    for lam, eta, mcw in [(1.0, 1.0, 1), (3.0, 0.4, 30)]:
        got = ctx.cv(genes(max_depth=1, eta=eta, min_child_weight=mcw, **{"lambda": lam}), orc.make_opts(nfold=4, num_boost_round=1))
        want = _stump_cv_by_hand(x, y, 4, lam, eta, mcw)
        assert abs(got["fitness"] - want) < 1e-6 * max(1.0, abs(want))
    ctx.close()


# ---- a second, structurally different restatement (recursive, float64, no histograms, no subtraction trick) for the
# parts of SURVEY.md A.5-A.9 that the stump test does not reach: several levels, alpha (ThresholdL1), max_delta_step
# (clamped weight and the matching gain), gamma as a pre-stop, min_child_weight on both sides, scale_pos_weight, the
# logistic objective and metric, base-score estimation, and several rounds of prediction updates.
def _sigmoid32(m):
    m = np.asarray(m, dtype=np.float32)
    return (np.float32(1.0) / (np.exp(np.minimum(-m, np.float32(88.7))) + np.float32(1.0) + np.float32(1e-16))).astype(np.float32)


def _grad(obj, margin, y, spw):
    w = np.where(y == np.float32(1.0), np.float32(spw), np.float32(1.0)).astype(np.float32)
    if obj == "binary:logistic":
        p = _sigmoid32(margin)
        return ((p - y) * w).astype(np.float32), (np.maximum(p * (np.float32(1.0) - p), np.float32(1e-16)) * w).astype(np.float32)
    return ((margin - y) * w).astype(np.float32), (np.ones_like(y) * w).astype(np.float32)


def _tl1(G, a):
    return G - a if G > a else G + a if G < -a else 0.0


def _weight(G, H, p):
    if H < p["min_child_weight"] or H <= 0:
        return 0.0
    w = -_tl1(G, p["alpha"]) / (H + p["lambda"])
    if p["max_delta_step"] != 0 and abs(w) > p["max_delta_step"]:
        w = np.copysign(p["max_delta_step"], w)
    return float(np.float32(w))


def _gain(G, H, p):
    if H <= 0:
        return 0.0
    if p["max_delta_step"] == 0:
        return _tl1(G, p["alpha"]) ** 2 / (H + p["lambda"])
    w = _weight(G, H, p)
    return -(2.0 * G * w + (H + p["lambda"]) * w * w)


def _grow(xs, g, h, rows, depth, p, out, path=()):
    """Returns nothing; fills `out` with (path -> leaf weight) and (path -> (feature, threshold)) entries."""
    G, H = float(g[rows].sum()), float(h[rows].sum())
    best = (0.0, None)
    if depth < p["max_depth"]:
        root = _gain(G, H, p)
        for j in range(xs.shape[1]):
            col = xs[rows, j]
            for thr in np.unique(col)[1:]:
                left = col < thr
                GL, HL = float(g[rows][left].sum()), float(h[rows][left].sum())
                if HL < p["min_child_weight"] or H - HL < p["min_child_weight"]:
                    continue
                chg = np.float32(np.float32(_gain(GL, HL, p)) + np.float32(_gain(G - GL, H - HL, p))) - np.float32(root)
                if chg > best[0]:
                    best = (float(chg), (j, thr))
    if best[1] is None or best[0] <= 1e-6 or best[0] < p["gamma"]:
        out[path] = _weight(G, H, p)
        return
    j, thr = best[1]
    out[path] = (j, thr)
    left = xs[rows, j] < thr
    _grow(xs, g, h, rows[left], depth + 1, p, out, path + (0,))
    _grow(xs, g, h, rows[~left], depth + 1, p, out, path + (1,))


def _apply(tree, xrow):
    path = ()
    while isinstance(tree[path], tuple):
        j, thr = tree[path]
        path += (0,) if xrow[j] < thr else (1,)
    return tree[path]


def _cv_by_hand(x, y, nfold, rounds, obj, metric, p):
    np.random.seed(0)
    chunks = np.array_split(np.random.permutation(len(y)), nfold)
    trace = np.zeros(rounds)
    for k in range(nfold):
        te = chunks[k]
        tr = np.concatenate([chunks[i] for i in range(nfold) if i != k])
        g0, h0 = _grad(obj, np.zeros(len(tr), dtype=np.float32), y[tr], p["scale_pos_weight"])
        stump = np.float32(-float(g0.astype(np.float64).sum()) / max(float(h0.astype(np.float64).sum()), 1e-6))
        base = np.float32(-np.log(np.float32(1.0) / _sigmoid32(stump) - np.float32(1.0))) if obj == "binary:logistic" else stump
        mtr = np.full(len(tr), base, dtype=np.float32)
        mte = np.full(len(te), base, dtype=np.float32)
        for r in range(rounds):
            g, h = _grad(obj, mtr, y[tr], p["scale_pos_weight"])
            tree = {}
            _grow(x[tr], g.astype(np.float64), h.astype(np.float64), np.arange(len(tr)), 0, p, tree)
            eta = np.float32(p["eta"])
            mtr = (mtr + np.array([eta * np.float32(_apply(tree, row)) for row in x[tr]], dtype=np.float32)).astype(np.float32)
            mte = (mte + np.array([eta * np.float32(_apply(tree, row)) for row in x[te]], dtype=np.float32)).astype(np.float32)
            if metric == "logloss":
                pr = _sigmoid32(mte).astype(np.float64)
                pr = np.clip(pr, 1e-16, 1 - 1e-16)
                e = -(y[te] * np.log(pr) + (1 - y[te]) * np.log(1 - pr))
                trace[r] += e.mean() / nfold
            else:
                d = (y[te] - mte).astype(np.float64)
                trace[r] += np.sqrt((d * d).mean()) / nfold
    return trace


@pytest.mark.parametrize("obj,metric,over", [
    ("reg:linear", "rmse", dict(max_depth=3, alpha=2.0, min_child_weight=12)),
    ("reg:linear", "rmse", dict(max_depth=3, max_delta_step=1, alpha=0.5, eta=0.7)),
    ("reg:linear", "rmse", dict(max_depth=4, gamma=6.0, **{"lambda": 0.5})),
    ("binary:logistic", "logloss", dict(max_depth=3, scale_pos_weight=2.5, min_child_weight=3)),
    ("binary:logistic", "logloss", dict(max_depth=2, max_delta_step=2, eta=1.0, scale_pos_weight=0.5)),
])
def test_multilevel_growth_matches_recursive_restatement(obj, metric, over):
    rng = np.random.default_rng(11)
    n = 360
    x = rng.integers(0, 7, size=(n, 4)).astype(np.float32)
    score = x[:, 0] - 0.8 * x[:, 1] + 1.5 * (x[:, 2] > 3) + rng.standard_normal(n) * 0.7
    if obj == "binary:logistic":
        y = (score > np.median(score)).astype(np.float32)
    else:
        y = score.astype(np.float32)
    g = genes(**over)
    rounds = 3
    ctx = orc.Context(x, y, 3)
    got = ctx.cv(g, orc.make_opts(objective=obj, eval_metric=metric, nfold=3, num_boost_round=rounds, early_stopping_rounds=0), trace=True)
    ctx.close()
    want = _cv_by_hand(x, y, 3, rounds, obj, metric, g)
    np.testing.assert_allclose(got["trace"], want, rtol=1e-6)


def test_training_reduces_validation_error(wine_ctx, wine):
    r = wine_ctx.cv(genes(eta=0.1), orc.make_opts(nfold=3, num_boost_round=30), trace=True)
    assert r["trace"][-1] < r["trace"][0] < float(np.std(wine[1].values)) * 1.01
    assert np.all(np.diff(r["trace"][:10]) < 0)


def test_sampling_and_edge_genes_run(wine_ctx):
    base = wine_ctx.cv(genes(max_depth=3), orc.make_opts(nfold=3, num_boost_round=5))["fitness"]
    for g in [genes(max_depth=3, subsample=0.5), genes(max_depth=3, colsample_bytree=0.3, colsample_bylevel=0.5),
              genes(max_depth=3, colsample_bytree=0.0002)]:
        v = wine_ctx.cv(g, orc.make_opts(nfold=3, num_boost_round=5))["fitness"]
        assert np.isfinite(v) and v != base
    # subsample = 0: every gradient is zero, the model never leaves the base score
    r = wine_ctx.cv(genes(subsample=0.0), orc.make_opts(nfold=3, num_boost_round=4), trace=True)
    assert np.all(r["trace"] == r["trace"][0])
    with pytest.raises(ValueError):
        wine_ctx.cv(genes(max_depth=11), orc.make_opts(nfold=3, num_boost_round=1))


def test_per_fit_path_equals_shared_context(wine):
    x, y = wine
    g = genes(max_depth=4)
    a = orc.cross_validate(x.values, y.values, g, nfold=3, num_boost_round=6)
    ctx = orc.Context(x.values, y.values, 3)
    b = ctx.cv(g, orc.make_opts(nfold=3, num_boost_round=6))["fitness"]
    ctx.close()
    assert a == b
