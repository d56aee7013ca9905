"""Individuals: genome, genes, reproduction, mutation and the lazy fitness cache.

Behavioural mirror of gentun/individuals.py:38-216 (Individual, XgboostIndividual).  Every use of
Python's module-level `random` happens in the same order and with the same arguments as there, so a
seeded run breeds bit-identical offspring.
"""
import math
import pprint
import random

from .cv_model import XgboostModel
from . import engine as _engine

_TINY = 1e-12


def random_log_uniform(minimum, maximum, base, eps=_TINY):
    """Draw uniformly on a log scale (base > 0: from minimum up; base < 0: mirrored from maximum
    down; base == 0: plain uniform).  One `random.uniform` call, like individuals.py:23-35."""
    if base == 0:
        return random.uniform(minimum, maximum)
    minimum += eps
    if base > 0:
        lo, hi = math.log(minimum, base), math.log(maximum, base)
        return base ** random.uniform(lo, hi)
    base = abs(base)
    lo, hi = math.log(eps, base), math.log(maximum - minimum, base)
    return maximum - base ** random.uniform(lo, hi)


def _draw_gene(spec):
    default, minimum, maximum, log_scale = spec
    if type(default) == int:
        return random.randint(minimum, maximum)
    return round(random_log_uniform(minimum, maximum, log_scale), 4)


class Individual(object):
    """Base species.  Subclasses supply a genome, a random gene generator and a fitness."""

    def __init__(self, x_train, y_train, genome, genes, crossover_rate, mutation_rate, additional_parameters=None):
        assert additional_parameters is None
        self.x_train, self.y_train = x_train, y_train
        self.genome = genome
        self.validate_genome()
        self.genes = genes
        self.validate_genes()
        self.crossover_rate, self.mutation_rate = crossover_rate, mutation_rate
        self.fitness = None

    def validate_genome(self):
        if type(self.genome) != dict:
            raise TypeError("Genome must be a dictionary.")
        if any(type(name) != str for name in self.genome):
            raise TypeError("Gene names must be strings.")

    def validate_genes(self):
        if set(self.genes.keys()) != set(self.genome.keys()):
            raise ValueError("Genes passed don't correspond to individual's genome.")

    def get_genes(self):
        return self.genes

    def get_genome(self):
        return self.genome

    @staticmethod
    def generate_random_genes(genome):
        raise NotImplementedError("Use a subclass with genes definition.")

    def evaluate_fitness(self):
        raise NotImplementedError("Use a subclass with genes definition.")

    def get_additional_parameters(self):
        raise NotImplementedError("Use a subclass with genes definition.")

    def get_fitness(self):
        if self.fitness is None:
            self.evaluate_fitness()
        return self.fitness

    def get_fitness_status(self):
        return self.fitness is not None

    def set_fitness(self, value):
        self.fitness = value

    def _spawn(self, genes):
        return self.__class__(self.x_train, self.y_train, self.genome, genes, self.crossover_rate,
                              self.mutation_rate, **self.get_additional_parameters())

    def reproduce(self, partner):
        """New individual; each gene comes from `partner` with probability crossover_rate."""
        assert self.__class__ == partner.__class__
        theirs = partner.get_genes()
        child = {name: (theirs[name] if random.random() < self.crossover_rate else mine)
                 for name, mine in self.get_genes().items()}
        return self._spawn(child)

    def crossover(self, partner):
        """In-place gene swap between the two parents."""
        assert self.__class__ == partner.__class__
        a, b = self.get_genes(), partner.get_genes()
        for name in a.keys():
            if random.random() < self.crossover_rate:
                a[name], b[name] = b[name], a[name]
                self.set_fitness(None)
                partner.set_fitness(None)

    def mutate(self):
        genes = self.get_genes()
        for name in list(genes.keys()):
            if random.random() < self.mutation_rate:
                genes[name] = _draw_gene(self.get_genome()[name])
                self.set_fitness(None)

    def copy(self):
        twin = self._spawn(self.genes.copy())
        twin.set_fitness(self.fitness)
        return twin

    def __str__(self):
        return pprint.pformat(self.genes)


class XgboostIndividual(Individual):
    """The 11-gene xgboost species of individuals.py:156-216."""

    def __init__(self, x_train, y_train, genome=None, genes=None, crossover_rate=0.5, mutation_rate=0.015,
                 booster='gbtree', objective='reg:linear', eval_metric='rmse', nfold=5,
                 num_boost_round=5000, early_stopping_rounds=100):
        if genome is None:
            genome = self.default_genome()
        if genes is None:
            genes = self.generate_random_genes(genome)
        super(XgboostIndividual, self).__init__(x_train, y_train, genome, genes, crossover_rate, mutation_rate)
        self.booster = booster
        self.objective = objective
        self.eval_metric = eval_metric
        self.nfold = nfold
        self.num_boost_round = num_boost_round
        self.early_stopping_rounds = early_stopping_rounds

    @staticmethod
    def default_genome():
        # name: (default, min, max, log-scale base) -- individuals.py:162-175
        return {
            'eta': (0.3, 0.001, 1.0, 10),
            'min_child_weight': (1, 0, 10, None),
            'max_depth': (6, 3, 10, None),
            'gamma': (0.0, 0.0, 10.0, 10),
            'max_delta_step': (0, 0, 10, None),
            'subsample': (1.0, 0.0, 1.0, -10),
            'colsample_bytree': (1.0, 0.0, 1.0, -10),
            'colsample_bylevel': (1.0, 0.0, 1.0, -10),
            'lambda': (1.0, 0.1, 10.0, 10),
            'alpha': (0.0, 0.0, 10.0, 10),
            'scale_pos_weight': (1.0, 0.0, 10.0, 0),
        }

    @staticmethod
    def generate_random_genes(genome):
        return {name: _draw_gene(spec) for name, spec in genome.items()}

    def get_additional_parameters(self):
        return {
            'booster': self.booster, 'objective': self.objective, 'eval_metric': self.eval_metric,
            'nfold': self.nfold, 'num_boost_round': self.num_boost_round,
            'early_stopping_rounds': self.early_stopping_rounds,
        }

    def evaluate_fitness(self):
        model = XgboostModel(self.x_train, self.y_train, self.genes, **self.get_additional_parameters())
        self.fitness = model.cross_validate()

    @classmethod
    def evaluate_batch(cls, individuals):
        """Batch hook used by Population: train every pending individual in one engine call
        (the shape of DistributedPopulation.evaluate_in_parallel, gentun/master.py:108-129)."""
        groups = {}
        for ind in individuals:
            params = ind.get_additional_parameters()
            key = (id(ind.x_train), id(ind.y_train), tuple(sorted(params.items())))
            groups.setdefault(key, (ind.x_train, ind.y_train, params, []))[3].append(ind)
        for x, y, params, members in groups.values():
            values = _engine.evaluate_many(x, y, [m.get_genes() for m in members], params)
            for m, v in zip(members, values):
                m.set_fitness(v)
