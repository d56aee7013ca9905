#!/usr/bin/env python
"""Generates tests/golden/ga_traces.json by running the UNMODIFIED reference (/root/reference/gentun)
with a stub fitness, so that gentun_b200's host-side GA can be checked for bit-identical random draws.

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_ga_golden.py
`pika` (hard import in gentun/__init__.py:4-5) is stubbed; xgboost is absent, so the species'
evaluate_fitness is replaced by a deterministic function of the genes.
"""
import contextlib
import hashlib
import io
import json
import math
import os
import random
import sys
import types

sys.modules.setdefault("pika", types.ModuleType("pika"))
sys.path.insert(0, "/root/reference")
with contextlib.redirect_stdout(io.StringIO()):
    import gentun  # noqa: E402
    from gentun.individuals import XgboostIndividual  # noqa: E402

SEED = 20240601


def stub_fitness(genes):
    return sum(math.sin(float(v) * (i + 1)) for i, (_, v) in enumerate(sorted(genes.items())))


class StubIndividual(XgboostIndividual):
    def evaluate_fitness(self):
        self.fitness = stub_fitness(self.genes)


def snapshot(pop):
    return [dict(pop[i].get_genes()) for i in range(pop.get_size())]


def key_orders(pop):
    return [list(pop[i].get_genes().keys()) for i in range(pop.get_size())]


def run_ga(pop, generations, cls=gentun.GeneticAlgorithm, **kw):
    ga = cls(pop, **kw)
    gens = []
    while ga.generation <= generations:
        ga.evolve_population()
        ga.generation += 1
        gens.append({"genes": snapshot(ga.population), "key_order": key_orders(ga.population)[:3],
                     "fitness_known": [ga.population[i].get_fitness_status() for i in range(ga.population.get_size())]})
    return gens


def main():
    out = {"seed": SEED, "python": sys.version.split()[0]}
    with contextlib.redirect_stdout(io.StringIO()):
        random.seed(SEED)
        pop = gentun.Population(StubIndividual, None, None, size=24, additional_parameters={"nfold": 3}, maximize=False)
        out["random_population"] = snapshot(pop)
        out["random_population_params"] = pop[0].get_additional_parameters()
        out["ga_minimize"] = run_ga(pop, 4)
        out["draw_after_ga_minimize"] = random.random()

        random.seed(SEED + 1)
        pop = gentun.Population(StubIndividual, None, None, size=16, maximize=True)
        out["ga_maximize_no_elitism"] = run_ga(pop, 3, tournament_size=3, elitism=False)
        out["draw_after_ga_maximize"] = random.random()

        random.seed(SEED + 2)
        grid = {"eta": [0.001, 0.005, 0.01, 0.015, 0.2], "max_depth": range(3, 11),
                "colsample_bytree": [0.80, 0.85, 0.90, 0.95, 1.0]}
        gpop = gentun.GridPopulation(StubIndividual, None, None, genes_grid=grid, additional_parameters={"nfold": 3}, maximize=False)
        out["grid_population"] = snapshot(gpop)
        out["grid_key_order"] = key_orders(gpop)[0]
        out["ga_grid"] = run_ga(gpop, 2)
        out["draw_after_ga_grid"] = random.random()

        random.seed(SEED + 3)
        pop = gentun.Population(StubIndividual, None, None, size=10, maximize=False)
        out["ga_roulette"] = run_ga(pop, 3, cls=gentun.RussianRouletteGA)
        out["draw_after_roulette"] = random.random()

        random.seed(SEED + 4)
        out["log_uniform"] = [gentun.individuals.random_log_uniform(a, b, c) for a, b, c in
                              [(0.001, 1.0, 10), (0.0, 10.0, 10), (0.0, 1.0, -10), (0.0, 10.0, 0), (0.1, 10.0, 10)] * 4]
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ga_traces.json")
    with open(path, "w") as fh:
        json.dump(out, fh, indent=0, sort_keys=True)
    print("wrote", path, hashlib.sha1(open(path, "rb").read()).hexdigest()[:12])


if __name__ == "__main__":
    main()
