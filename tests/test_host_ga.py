"""Host-side GA logic against traces recorded from the unmodified reference
(tests/golden/make_ga_golden.py): seeded offspring must be bit-identical."""
import contextlib
import io
import json
import math
import os
import random

import pytest

import gentun_b200 as gentun
from gentun_b200.genetics import XgboostIndividual

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "ga_traces.json")))
SEED = GOLD["seed"]


def stub_fitness(genes):
    return sum(math.sin(float(v) * (i + 1)) for i, (_, v) in enumerate(sorted(genes.items())))


class StubIndividual(XgboostIndividual):
    def evaluate_fitness(self):
        self.fitness = stub_fitness(self.genes)


def snapshot(pop):
    return [dict(pop[i].get_genes()) for i in range(pop.get_size())]


def run_ga(pop, generations, cls=gentun.GeneticAlgorithm, **kw):
    ga = cls(pop, **kw)
    gens = []
    while ga.generation <= generations:
        ga.evolve_population()
        ga.generation += 1
        gens.append({"genes": snapshot(ga.population),
                     "key_order": [list(ga.population[i].get_genes().keys()) for i in range(3)],
                     "fitness_known": [ga.population[i].get_fitness_status() for i in range(ga.population.get_size())]})
    return gens


def check_gens(got, want):
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g["genes"] == w["genes"]
        assert g["key_order"] == w["key_order"]
        assert g["fitness_known"] == w["fitness_known"]


@pytest.fixture(autouse=True)
def quiet():
    with contextlib.redirect_stdout(io.StringIO()):
        yield


def test_random_population_and_ga_minimize():
    random.seed(SEED)
    pop = gentun.Population(StubIndividual, None, None, size=24, additional_parameters={"nfold": 3}, maximize=False)
    assert snapshot(pop) == GOLD["random_population"]
    assert pop[0].get_additional_parameters() == GOLD["random_population_params"]
    check_gens(run_ga(pop, 4), GOLD["ga_minimize"])
    assert random.random() == GOLD["draw_after_ga_minimize"]


def test_tournaments_build_no_throwaway_populations():
    """the reference constructs a population per tournament, 2(P - 1) per generation
    (algorithms.py:50-56); here one generation constructs exactly one (the next generation), with the same offspring."""
    built = []

    class Counting(gentun.Population):
        def __init__(self, *a, **kw):
            built.append(1)
            super(Counting, self).__init__(*a, **kw)

    random.seed(SEED)
    pop = Counting(StubIndividual, None, None, size=24, additional_parameters={"nfold": 3}, maximize=False)
    del built[:]
    gens = run_ga(pop, 4)
    assert len(built) == 4
    check_gens(gens, GOLD["ga_minimize"])
    # a contender whose fitness is unknown still reaches the batch point through a population
    ga = gentun.GeneticAlgorithm(Counting(StubIndividual, None, None, size=8, maximize=False), tournament_size=8)
    del built[:]
    winner = ga.tournament_select()
    assert len(built) == 1 and winner.get_fitness() == min(ga.population[i].get_fitness() for i in range(8))


def test_ga_maximize_without_elitism():
    random.seed(SEED + 1)
    pop = gentun.Population(StubIndividual, None, None, size=16, maximize=True)
    check_gens(run_ga(pop, 3, tournament_size=3, elitism=False), GOLD["ga_maximize_no_elitism"])
    assert random.random() == GOLD["draw_after_ga_maximize"]


def test_grid_population_and_ga():
    random.seed(SEED + 2)
    grid = {"eta": [0.001, 0.005, 0.01, 0.015, 0.2], "max_depth": range(3, 11),
            "colsample_bytree": [0.80, 0.85, 0.90, 0.95, 1.0]}
    gpop = gentun.GridPopulation(StubIndividual, None, None, genes_grid=grid, additional_parameters={"nfold": 3}, maximize=False)
    assert gpop.get_size() == 200
    assert snapshot(gpop) == GOLD["grid_population"]
    assert list(gpop[0].get_genes().keys()) == GOLD["grid_key_order"]
    check_gens(run_ga(gpop, 2), GOLD["ga_grid"])
    assert random.random() == GOLD["draw_after_ga_grid"]


def test_russian_roulette():
    random.seed(SEED + 3)
    pop = gentun.Population(StubIndividual, None, None, size=10, maximize=False)
    check_gens(run_ga(pop, 3, cls=gentun.RussianRouletteGA), GOLD["ga_roulette"])
    assert random.random() == GOLD["draw_after_roulette"]


def test_random_log_uniform():
    random.seed(SEED + 4)
    got = [gentun.random_log_uniform(a, b, c) for a, b, c in
           [(0.001, 1.0, 10), (0.0, 10.0, 10), (0.0, 1.0, -10), (0.0, 10.0, 0), (0.1, 10.0, 10)] * 4]
    assert got == GOLD["log_uniform"]


def test_elite_is_same_object_and_fitness_kept():
    random.seed(7)
    pop = gentun.Population(StubIndividual, None, None, size=8, maximize=False)
    ga = gentun.GeneticAlgorithm(pop)
    best = pop.get_fittest()
    ga.evolve_population()
    assert ga.population[0] is best and best.get_fitness_status()
    # run(G) evaluates G generations; the population bred last is left unevaluated
    assert not all(ga.population[i].get_fitness_status() for i in range(1, 8))


def test_population_validation_and_protocol():
    with pytest.raises(ValueError):
        gentun.Population(StubIndividual, None, None)
    with pytest.raises(ValueError):
        gentun.GridPopulation(StubIndividual, None, None)
    with pytest.raises(ValueError):
        gentun.GridPopulation(StubIndividual, None, None, genes_grid={"nope": [1]})
    with pytest.raises(ValueError):
        StubIndividual(None, None, genes={"eta": 0.1})
    with pytest.raises(TypeError):
        gentun.Individual(None, None, [("eta", 1)], {}, 0.5, 0.1)
    ind = StubIndividual(None, None, nfold=3, num_boost_round=10)
    twin = ind.copy()
    assert twin.get_genes() == ind.get_genes() and twin.get_genes() is not ind.get_genes()
    assert twin.get_additional_parameters() == ind.get_additional_parameters()
    assert ind.get_additional_parameters()["nfold"] == 3
    # mutation resets the cached fitness
    ind.get_fitness()
    ind.mutation_rate = 1.0
    ind.mutate()
    assert not ind.get_fitness_status()


def test_batch_hook_is_skipped_for_overridden_species():
    from gentun_b200.population import _batch_hook
    assert _batch_hook(StubIndividual) is None
    assert _batch_hook(XgboostIndividual) is not None


def test_batch_hook_groups_and_memo(monkeypatch):
    """Population.get_fittest sends all pending individuals to the engine in one call and reuses
    fitness of identical gene sets."""
    from gentun_b200 import engine as eng_mod
    calls = []

    class FakeEngine(object):
        def evaluate(self, genes_list, **params):
            calls.append((len(genes_list), params["nfold"]))
            import numpy as np
            return np.array([stub_fitness({k: g[k] for k in eng_mod.GENE_ORDER}) for g in genes_list]), None

    fake = FakeEngine()
    monkeypatch.setattr(eng_mod, "engine_for", lambda x, y, device=None: fake)
    eng_mod._memo.clear()
    random.seed(3)
    x, y = object(), object()
    pop = gentun.Population(XgboostIndividual, x, y, size=6, additional_parameters={"nfold": 3}, maximize=False)
    dup = pop[0].copy()
    dup.set_fitness(None)
    pop.add_individual(dup)
    best = pop.get_fittest()
    assert calls == [(6, 3)]          # 7 pending, one duplicate trained once
    assert all(pop[i].get_fitness_status() for i in range(7))
    assert best.get_fitness() == min(pop[i].get_fitness() for i in range(7))
    pop.get_fittest()
    assert len(calls) == 1            # nothing pending: no engine call
    model = gentun.XgboostModel(x, y, pop[1].get_genes(), nfold=3)
    assert model.cross_validate() == pop[1].get_fitness() and len(calls) == 1   # memo hit
    eng_mod._memo.clear()


def test_engine_cache_follows_the_data(monkeypatch):
    """Engines and memoised fitness are tied to the data: a new dataset never sees another dataset's values (not even
    at a recycled address), data changed in place gets a new engine, and an engine is closed when its data die."""
    import gc
    import numpy as np
    from gentun_b200 import engine as eng_mod
    made, closed = [], []

    class FakeEngine(object):
        def __init__(self, x, y, device=0):
            self.tag = float(np.asarray(x)[0, 0])
            made.append(self.tag)

        def evaluate(self, genes_list, **params):
            return np.array([self.tag + g["eta"] for g in genes_list]), None

        def close(self):
            closed.append(self.tag)

    monkeypatch.setattr(eng_mod, "Engine", FakeEngine)
    eng_mod.clear_engines()
    g = dict(eng_mod.XGB_DEFAULTS)
    params = {"nfold": 3}
    x1, y1 = np.full((8, 2), 1.0), np.zeros(8)
    assert eng_mod.evaluate_many(x1, y1, [g], params) == [1.3]
    assert eng_mod.evaluate_many(x1, y1, [g], params) == [1.3] and made == [1.0]      # memo + cached engine
    x1[:] = 5.0                                                                          # rescaled in place
    assert eng_mod.evaluate_many(x1, y1, [g], params) == [5.3] and made == [1.0, 5.0] and closed == [1.0]
    del x1
    gc.collect()
    assert closed == [1.0, 5.0] and not eng_mod._engines and not eng_mod._memo         # data died: engine and memo gone
    for v in (7.0, 9.0):                                                                 # datasets in a loop, ids may recycle
        xv = np.full((8, 2), v)
        assert eng_mod.evaluate_many(xv, y1, [g], params) == [v + 0.3]
        del xv
        gc.collect()
    assert closed == [1.0, 5.0, 7.0, 9.0]
    eng_mod.clear_engines()


def test_partial_genomes_take_xgboost_defaults():
    """A genome that tunes two genes trains with xgboost's defaults for the rest (XgboostModel dict.update()s only what
    it is given, xgboost_models.py:16-22); a parameter this engine cannot honour is refused, not ignored."""
    from gentun_b200.engine import genes_matrix, GENE_ORDER
    m = genes_matrix([{"eta": 0.1, "max_depth": 4, "booster": "gbtree", "silent": 1}])
    want = dict(eta=0.1, min_child_weight=1, max_depth=4, gamma=0, max_delta_step=0, subsample=1, colsample_bytree=1,
                colsample_bylevel=1, alpha=0, scale_pos_weight=1)
    want["lambda"] = 1
    assert [m[0, j] for j in range(11)] == [float(want[k]) for k in GENE_ORDER]
    with pytest.raises(NotImplementedError):
        genes_matrix([{"eta": 0.1, "tree_method": "exact"}])
    with pytest.raises(ValueError):
        genes_matrix([{"eta": 0.1, "not_a_parameter": 3}])


def test_unsupported_parameters_raise():
    from gentun_b200.engine import fit_opts
    with pytest.raises(NotImplementedError):
        fit_opts(booster="gblinear")
    with pytest.raises(NotImplementedError):
        fit_opts(objective="multi:softmax")
    with pytest.raises(NotImplementedError):
        fit_opts(eval_metric="auc")
