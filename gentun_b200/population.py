"""Populations: the batch-evaluation point of the GA (mirror of gentun/populations.py:10-105)."""
import itertools


def _batch_hook(species):
    """species.evaluate_batch, unless a subclass replaced evaluate_fitness below the class that
    defines the hook (then the subclass's own per-individual fitness must be honoured)."""
    hook = getattr(species, 'evaluate_batch', None)
    if hook is None:
        return None
    for klass in species.__mro__:
        if 'evaluate_fitness' in klass.__dict__ or 'evaluate_batch' in klass.__dict__:
            return hook if 'evaluate_batch' in klass.__dict__ else None
    return None


class Population(object):
    """Individuals of one species.  get_fittest() first evaluates, in one batch, every individual
    whose fitness is unknown, then returns the max/min by fitness (first wins ties)."""

    def __init__(self, species, x_train, y_train, individual_list=None, size=None,
                 crossover_rate=0.5, mutation_rate=0.015, maximize=True, additional_parameters=None):
        self.x_train, self.y_train = x_train, y_train
        self.species = species
        self.maximize = maximize
        if individual_list is None and size is None:
            raise ValueError("Either pass a list of individuals or a population size for a random population.")
        if individual_list is None:
            extra = additional_parameters or {}
            self.individuals = [
                species(x_train, y_train, crossover_rate=crossover_rate, mutation_rate=mutation_rate, **extra)
                for _ in range(size)
            ]
            self.population_size = size
            print("Initializing a random population. Size: {}".format(size))
        else:
            assert all(type(individual) is species for individual in individual_list)
            self.individuals = individual_list
            self.population_size = len(individual_list)

    def add_individual(self, individual):
        assert type(individual) is self.species
        self.individuals.append(individual)
        self.population_size += 1

    def get_species(self):
        return self.species

    def get_size(self):
        return self.population_size

    def get_data(self):
        return self.x_train, self.y_train

    def get_fitness_criteria(self):
        return self.maximize

    def __getitem__(self, item):
        return self.individuals[item]

    def evaluate_pending(self):
        pending = [ind for ind in self.individuals if not ind.get_fitness_status()]
        if not pending:
            return 0
        hook = _batch_hook(self.species)
        if hook is not None:
            hook(pending)
        else:
            for ind in pending:
                ind.get_fitness()
        return len(pending)

    def get_fittest(self):
        self.evaluate_pending()
        pick = max if self.maximize else min
        return pick(self.individuals, key=lambda ind: ind.get_fitness())


class GridPopulation(Population):
    """Population built from the cartesian product of a gene grid (grid keys first, then the
    remaining genome keys at their defaults)."""

    def __init__(self, species, x_train, y_train, individual_list=None, genes_grid=None,
                 crossover_rate=0.5, mutation_rate=0.015, maximize=True, additional_parameters=None):
        if individual_list is None and genes_grid is None:
            raise ValueError("Either pass a list of individuals or a grid definition.")
        if genes_grid is not None:
            genome = species(None, None).get_genome()   # draws one random gene set, as the reference does
            if not set(genes_grid.keys()).issubset(set(genome.keys())):
                raise ValueError("Some grid parameters do not belong to the species' genome")
            axes = dict(genes_grid)
            for gene, spec in genome.items():
                axes.setdefault(gene, [spec[0]])
            names = list(axes.keys())
            extra = additional_parameters or {}
            individual_list = [
                species(x_train, y_train, genes=dict(zip(names, combo)), crossover_rate=crossover_rate,
                        mutation_rate=mutation_rate, **extra)
                for combo in itertools.product(*axes.values())
            ]
            print("Initializing a grid population. Size: {}".format(len(individual_list)))
        super(GridPopulation, self).__init__(species, x_train, y_train, individual_list, None, crossover_rate,
                                             mutation_rate, maximize, additional_parameters)
