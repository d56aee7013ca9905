"""Genetic-algorithm drivers.

Behavioural mirror of gentun/algorithms.py:9-99: the same console lines, the same use of Python's module-level
`random` in the same order (Appendix B of SURVEY.md), so a seeded run breeds the same offspring.  Only
`population.get_fittest()` reaches the GPU; selection, crossover and mutation are host Python.
"""
import random

_BANNER = "Starting genetic algorithm...\n"


def _announce(generation, fittest):
    """The three progress lines of the reference (fitness rounded to 4 decimals)."""
    print("Evaluating generation #{}...".format(generation))
    print("Fittest individual is:")
    print(fittest)
    print("Fitness value is: {}\n".format(round(fittest.get_fitness(), 4)))


class GeneticAlgorithm(object):
    """Generational GA: optional elitism, tournament selection, uniform crossover, per-gene mutation."""

    def __init__(self, population, tournament_size=5, elitism=True):
        self.population = population
        self.x_train, self.y_train = population.get_data()
        self.tournament_size = tournament_size
        self.elitism = elitism
        self.generation = 1

    def get_population_type(self):
        return self.population.__class__

    def _spawn_population(self, members):
        """A population of the same class around `members`.  The reference re-creates the population class with
        exactly these arguments for every generation (and for every tournament), which is why engines and broker
        credentials are kept outside the instance."""
        source = self.population
        return self.get_population_type()(source.get_species(), self.x_train, self.y_train, individual_list=members,
                                          maximize=source.get_fitness_criteria())

    def run(self, max_generations):
        print(_BANNER)
        while self.generation <= max_generations:
            self.evolve_population()
            self.generation += 1

    def tournament_select(self):
        """Fittest of `tournament_size` members drawn without replacement (algorithms.py:50-56: same `random.sample`
        call, same first-wins tie rule).  The reference wraps the contenders in a throw-away population, 2(P - 1) of them
        per generation, only to call get_fittest on it; every member already has its fitness here (evolve_population
        evaluated the generation in one batch), so the winner is picked directly.  Contenders without a fitness -- a
        caller using this method on its own -- still go through a population, i.e. through the batch point."""
        size = self.population.get_size()
        contenders = [self.population[i] for i in random.sample(range(size), self.tournament_size)]
        if all(ind.get_fitness_status() for ind in contenders):
            pick = max if self.population.get_fitness_criteria() else min
            return pick(contenders, key=lambda ind: ind.get_fitness())
        return self._spawn_population(contenders).get_fittest()

    def evolve_population(self):
        _announce(self.generation, self.population.get_fittest())
        next_generation = self._spawn_population([])
        if self.elitism:
            next_generation.add_individual(self.population.get_fittest())   # the same object: its fitness is kept
        wanted = self.population.get_size()
        while next_generation.get_size() < wanted:
            mother = self.tournament_select()
            father = self.tournament_select()
            child = mother.reproduce(father)
            child.mutate()
            next_generation.add_individual(child)
        self.population = next_generation


def _roulette_weights(scores, maximize, eps):
    """Selection weights of the Genetic-CNN paper: fitness (or its inverse) shifted so the worst gets zero."""
    raw = list(scores) if maximize else [1 / (score + eps) for score in scores]
    floor = min(raw)
    shifted = [value - floor for value in raw]
    if sum(shifted) == .0:
        shifted = [1. for _ in shifted]
    return shifted


class RussianRouletteGA(GeneticAlgorithm):
    """Fitness-proportional resampling, then pairwise crossover or mutation."""

    def __init__(self, population, crossover_probability=0.2, mutation_probability=0.8):
        super(RussianRouletteGA, self).__init__(population)
        self.crossover_probability = crossover_probability
        self.mutation_probability = mutation_probability

    def evolve_population(self, eps=1e-15):
        _announce(self.generation, self.population.get_fittest())
        current = self.population
        size = current.get_size()
        weights = _roulette_weights([current[i].get_fitness() for i in range(size)], current.get_fitness_criteria(), eps)
        picks = random.choices(range(size), weights=weights, k=size)
        resampled = self._spawn_population([current[i].copy() for i in picks])
        for i in range(resampled.get_size() // 2):
            first, second = resampled[i], resampled[i + 1]
            if random.random() < self.crossover_probability:
                first.crossover(second)
                continue
            if random.random() < self.mutation_probability:
                first.mutate()
            if random.random() < self.mutation_probability:
                second.mutate()
        self.population = resampled
