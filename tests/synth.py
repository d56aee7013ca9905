"""Synthetic data of SURVEY.md 8(d): shared by the tests and bench.py."""
import numpy as np


def make_dataset(n, f, seed, classification=False):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, f), dtype=np.float32)
    for j in range(f):
        if j % 10 == 9:
            x[:, j] = rng.integers(0, 12, n).astype(np.float32)
        elif j % 10 == 8:
            x[:, j] = np.exp(x[:, j])
    k = min(8, f)
    wv = rng.standard_normal(k).astype(np.float32)
    score = x[:, :k] @ wv
    if f > 1:
        score = score + 0.5 * x[:, 0] * x[:, 1]
    if f > 2:
        score = score + np.sin(x[:, 2])
    if classification:
        z = score - np.quantile(score, 0.7)
        p = 1.0 / (1.0 + np.exp(-z))
        y = (rng.random(n) < p).astype(np.float32)
    else:
        y = (score + 0.5 * rng.standard_normal(n)).astype(np.float32)
    return np.ascontiguousarray(x, dtype=np.float32), np.ascontiguousarray(y, dtype=np.float32)


def seeded_population_genes(size, seed=20240601):
    """Gene dicts of a random Population, drawn exactly like XgboostIndividual.generate_random_genes."""
    import contextlib
    import io
    import random
    from gentun_b200 import Population, XgboostIndividual
    state = random.getstate()
    random.seed(seed)
    with contextlib.redirect_stdout(io.StringIO()):
        pop = Population(XgboostIndividual, None, None, size=size)
    random.setstate(state)
    return [dict(pop[i].get_genes()) for i in range(size)]
