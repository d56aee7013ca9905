#!/usr/bin/env python
"""BASELINE.json config 2 on the drop-in API: grid search over eta x max_depth x colsample_bytree
(5 x 8 x 5 = 200 individuals) on the white-wine data, 3-fold CV, lower rmse is better.

    python examples/grid_wine_quality.py [generations]     # 1 generation = plain grid search
"""
import argparse
import os
import random
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

import pandas as pd  # noqa: E402

import gentun_b200 as gentun  # noqa: E402

GRID = {
    'eta': [0.001, 0.005, 0.01, 0.015, 0.2],
    'max_depth': range(3, 11),
    'colsample_bytree': [0.80, 0.85, 0.90, 0.95, 1.0],
}


def load_wine():
    table = pd.read_csv(os.path.join(os.path.dirname(HERE), 'tests', 'data', 'winequality-white.csv'), delimiter=';')
    return table.drop(['quality'], axis=1), table['quality']


def main():
    parser = argparse.ArgumentParser(description=__doc__)
    parser.add_argument('generations', nargs='?', type=int, default=1)
    args = parser.parse_args()
    random.seed(20240601)
    features, target = load_wine()
    started = time.time()
    population = gentun.GridPopulation(gentun.XgboostIndividual, features, target, genes_grid=dict(GRID),
                                       additional_parameters={'nfold': 3}, maximize=False)
    gentun.GeneticAlgorithm(population).run(args.generations)
    print("elapsed: %.1f s" % (time.time() - started))


if __name__ == '__main__':
    main()
