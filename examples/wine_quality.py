#!/usr/bin/env python
"""BASELINE.json config 1: the reference's tests/test_wine-quality.py with `import gentun_b200 as gentun`.
Population of 100 random XgboostIndividuals, nfold=3, minimise rmse, 10 generations (991 CV fits)."""
import os
import random
import sys
import time

import pandas as pd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

if __name__ == '__main__':
    from gentun_b200 import GeneticAlgorithm, Population, XgboostIndividual

    generations = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    size = int(sys.argv[2]) if len(sys.argv) > 2 else 100
    random.seed(20240601)
    data = pd.read_csv(os.path.join(ROOT, 'tests', 'data', 'winequality-white.csv'), delimiter=';')
    y_train = data['quality']
    x_train = data.drop(['quality'], axis=1)
    t0 = time.time()
    pop = Population(
        XgboostIndividual, x_train, y_train, size=size,
        additional_parameters={'nfold': 3}, maximize=False
    )
    ga = GeneticAlgorithm(pop)
    ga.run(generations)
    print("elapsed: %.1f s for %d generations of %d individuals" % (time.time() - t0, generations, size))
