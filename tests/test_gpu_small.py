"""The shared-memory-resident kernel (csrc/small.cu: one persistent CTA per (individual, fold) job, one launch per
evaluation) against the streaming kernels on the same data: both regimes must give the same bits, and through them the
same parity with the oracle.  Data: the reference's own winequality-white.csv (configs C1 / C2) and small synthetic sets."""
import numpy as np
import pytest

from oracle import orc
from tests.synth import make_dataset, seeded_population_genes
from tests.test_gpu_parity import GENE_SETS, genes, rel, REL_TOL

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wine_np(wine):
    x, y = wine
    return np.ascontiguousarray(x.values, dtype=np.float32), np.ascontiguousarray(y.values, dtype=np.float32)


@pytest.fixture(scope="module")
def wine_engines(wine_np):
    from gentun_b200.engine import Engine
    x, y = wine_np
    small, big = Engine(x, y), Engine(x, y, streaming_only=True)
    yield small, big
    small.close()
    big.close()


def test_small_regime_is_one_launch_and_bit_equal_to_streaming(wine_engines):
    small, big = wine_engines
    gs = GENE_SETS + [genes(subsample=0.7, max_depth=6), genes(colsample_bytree=0.5, colsample_bylevel=0.6, max_depth=7),
                      genes(max_depth=10, eta=0.02, min_child_weight=0, gamma=0.3)]
    a, ra, ea = small.evaluate(gs, trace=True, nfold=3, num_boost_round=30)
    assert small.stats()["kernel_launches"] == 1
    b, rb, eb = big.evaluate(gs, trace=True, nfold=3, num_boost_round=30)
    assert big.stats()["kernel_launches"] > 100
    assert np.array_equal(ra, rb)
    assert np.array_equal(ea["trace"], eb["trace"])          # every round's CV mean, bit for bit
    assert np.array_equal(a, b)


def test_small_regime_early_stopping_and_seeded_population(wine_engines, wine_np):
    small, big = wine_engines
    gs = seeded_population_genes(24) + [genes(eta=0.9, max_depth=8), genes(eta=0.5)]
    params = dict(nfold=3, num_boost_round=300, early_stopping_rounds=12)
    a, ra = small.evaluate(gs, **params)
    b, rb = big.evaluate(gs, **params)
    assert np.array_equal(ra, rb) and np.array_equal(a, b)
    assert ra.min() < 300                                     # somebody stopped early
    x, y = wine_np
    ctx = orc.Context(x, y, 3)
    for i in (0, 24, 25):
        want = ctx.cv(gs[i], orc.make_opts(**params))
        assert ra[i] == want["rounds"] and rel(a[i], want["fitness"]) < REL_TOL
    ctx.close()


@pytest.mark.parametrize("n,f,nfold,cls", [(2500, 7, 5, False), (4000, 40, 4, True), (130, 3, 2, False), (300, 1, 2, False), (8190, 5, 2, False)])
def test_small_regime_other_shapes(n, f, nfold, cls):
    """Other fold counts, more than 32 features (two mask words), logistic / logloss, a tiny set, a single feature, the largest train slice the regime takes (4095 rows = 32 chunks)."""
    from gentun_b200.engine import Engine
    x, y = make_dataset(n, f, seed=40 + f, classification=cls)
    small, big = Engine(x, y), Engine(x, y, streaming_only=True)
    gs = [genes(max_depth=6, min_child_weight=2), genes(max_depth=9, eta=0.1, min_child_weight=0, scale_pos_weight=1.8 if cls else 1.0),
          genes(max_depth=4, subsample=0.8, colsample_bytree=0.7, max_delta_step=2, alpha=0.4)]
    params = dict(nfold=nfold, num_boost_round=20, early_stopping_rounds=5)
    if cls:
        params.update(objective="binary:logistic", eval_metric="logloss")
    a, ra = small.evaluate(gs, **params)
    assert small.stats()["kernel_launches"] == 1
    b, rb = big.evaluate(gs, **params)
    assert np.array_equal(ra, rb) and np.array_equal(a, b)
    ctx = orc.Context(x, y, nfold)
    want = ctx.cv(gs[0], orc.make_opts(**params))
    assert ra[0] == want["rounds"] and rel(a[0], want["fitness"]) < REL_TOL
    ctx.close()
    small.close()
    big.close()


def test_small_regime_many_individuals_queue(wine_engines):
    """More individuals than resident teams: the teams take them from the queue; order of submission does not matter."""
    small, _ = wine_engines
    gs = seeded_population_genes(150)
    a, ra = small.evaluate(gs, nfold=3, num_boost_round=6)
    b, rb = small.evaluate(gs[::-1], nfold=3, num_boost_round=6)
    assert np.array_equal(a, b[::-1]) and np.array_equal(ra, rb[::-1])
