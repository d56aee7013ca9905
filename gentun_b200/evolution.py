"""Genetic-algorithm drivers (mirror of gentun/algorithms.py:9-99).  Selection, crossover and
mutation stay in host Python; only get_fittest() reaches the GPU."""
import random


class GeneticAlgorithm(object):
    """Tournament selection with optional elitism."""

    def __init__(self, population, tournament_size=5, elitism=True):
        self.population = population
        self.x_train, self.y_train = population.get_data()
        self.tournament_size = tournament_size
        self.elitism = elitism
        self.generation = 1

    def get_population_type(self):
        return self.population.__class__

    def _empty_like(self, members):
        return self.get_population_type()(self.population.get_species(), self.x_train, self.y_train,
                                          individual_list=members, maximize=self.population.get_fitness_criteria())

    def run(self, max_generations):
        print("Starting genetic algorithm...\n")
        while self.generation <= max_generations:
            self.evolve_population()
            self.generation += 1

    def _report(self):
        print("Evaluating generation #{}...".format(self.generation))
        fittest = self.population.get_fittest()
        print("Fittest individual is:")
        print(fittest)
        print("Fitness value is: {}\n".format(round(fittest.get_fitness(), 4)))
        return fittest

    def evolve_population(self):
        self._report()
        offspring = self._empty_like([])
        if self.elitism:
            offspring.add_individual(self.population.get_fittest())
        target = self.population.get_size()
        while offspring.get_size() < target:
            child = self.tournament_select().reproduce(self.tournament_select())
            child.mutate()
            offspring.add_individual(child)
        self.population = offspring

    def tournament_select(self):
        drawn = random.sample(range(self.population.get_size()), self.tournament_size)
        return self._empty_like([self.population[i] for i in drawn]).get_fittest()


class RussianRouletteGA(GeneticAlgorithm):
    """Fitness-proportional selection of the Genetic CNN paper."""

    def __init__(self, population, crossover_probability=0.2, mutation_probability=0.8):
        super(RussianRouletteGA, self).__init__(population)
        self.crossover_probability = crossover_probability
        self.mutation_probability = mutation_probability

    def evolve_population(self, eps=1e-15):
        self._report()
        size = self.population.get_size()
        scores = [self.population[i].get_fitness() for i in range(size)]
        weights = scores if self.population.get_fitness_criteria() else [1 / (s + eps) for s in scores]
        floor = min(weights)
        weights = [wgt - floor for wgt in weights]
        if sum(weights) == .0:
            weights = [1. for _ in range(size)]
        chosen = random.choices(range(size), weights=weights, k=size)
        offspring = self._empty_like([self.population[i].copy() for i in chosen])
        for i in range(offspring.get_size() // 2):
            if random.random() < self.crossover_probability:
                offspring[i].crossover(offspring[i + 1])
            else:
                if random.random() < self.mutation_probability:
                    offspring[i].mutate()
                if random.random() < self.mutation_probability:
                    offspring[i + 1].mutate()
        self.population = offspring
