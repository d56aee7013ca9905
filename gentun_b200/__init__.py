"""gentun_b200: gentun's public names (gentun/__init__.py:2-11) over a B200-native fitness engine.

Drop-in use:  `import gentun_b200 as gentun`, then Population / GridPopulation / GeneticAlgorithm /
XgboostIndividual / XgboostModel exactly as in the reference's tests/test_wine-quality.py.
"""
from .evolution import GeneticAlgorithm, RussianRouletteGA
from .population import Population, GridPopulation
from .genetics import Individual, XgboostIndividual, random_log_uniform
from .cv_model import GentunModel, XgboostModel
from .engine import Engine, EngineError, clear_engines
from .distributed import DistributedPopulation, DistributedGridPopulation, GentunWorker

__all__ = ["GeneticAlgorithm", "RussianRouletteGA", "Population", "GridPopulation", "Individual",
           "XgboostIndividual", "GentunModel", "XgboostModel", "Engine", "EngineError", "clear_engines",
           "random_log_uniform", "DistributedPopulation", "DistributedGridPopulation", "GentunWorker"]
__version__ = "0.1.0"
