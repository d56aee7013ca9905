#!/usr/bin/env python
"""BASELINE.json config 2: the reference's tests/test_grid_wine-quality.py grid (5 x 8 x 5 = 200 individuals)
evaluated as one generation (grid search)."""
import os
import random
import sys
import time

import pandas as pd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

if __name__ == '__main__':
    from gentun_b200 import GeneticAlgorithm, GridPopulation, XgboostIndividual

    random.seed(20240601)
    data = pd.read_csv(os.path.join(ROOT, 'tests', 'data', 'winequality-white.csv'), delimiter=';')
    y_train = data['quality']
    x_train = data.drop(['quality'], axis=1)
    grid = {
        'eta': [0.001, 0.005, 0.01, 0.015, 0.2],
        'max_depth': range(3, 11),
        'colsample_bytree': [0.80, 0.85, 0.90, 0.95, 1.0],
    }
    t0 = time.time()
    pop = GridPopulation(
        XgboostIndividual, x_train, y_train, genes_grid=grid,
        additional_parameters={'nfold': 3}, maximize=False
    )
    ga = GeneticAlgorithm(pop)
    ga.run(int(sys.argv[1]) if len(sys.argv) > 1 else 1)
    print("elapsed: %.1f s" % (time.time() - t0))
