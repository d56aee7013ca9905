"""Times one generation of BASELINE config 1 and the grid of config 2 (winequality-white, nfold 3, <= 5000 rounds, early stop 100)."""
import os, sys, time, random, contextlib, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pandas as pd
import gentun_b200 as gentun

d = pd.read_csv('tests/data/winequality-white.csv', delimiter=';')
y = d['quality']; x = d.drop(['quality'], axis=1)
random.seed(20240601)
with contextlib.redirect_stdout(io.StringIO()):
    pop = gentun.Population(gentun.XgboostIndividual, x, y, size=100, additional_parameters={'nfold': 3}, maximize=False)
t = time.time(); best = pop.get_fittest(); dt = time.time() - t
print("C1: one generation (100 individuals): %.2f s, best rmse %.4f" % (dt, best.get_fitness()), flush=True)
grid = {'eta': [0.001, 0.005, 0.01, 0.015, 0.2], 'max_depth': range(3, 11), 'colsample_bytree': [0.80, 0.85, 0.90, 0.95, 1.0]}
with contextlib.redirect_stdout(io.StringIO()):
    gp = gentun.GridPopulation(gentun.XgboostIndividual, x, y, genes_grid=grid, additional_parameters={'nfold': 3}, maximize=False)
t = time.time(); best = gp.get_fittest(); dt = time.time() - t
print("C2: grid of %d individuals: %.2f s, best rmse %.4f" % (gp.get_size(), dt, best.get_fitness()), flush=True)
