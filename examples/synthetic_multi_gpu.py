#!/usr/bin/env python
"""BASELINE.json config 3 through the drop-in API: XgboostIndividual regressor on synthetic 1M x 50 data,
Population of 256, nfold=5, 100 boosting rounds.  Single GPU:  python examples/synthetic_multi_gpu.py
All GPUs of the box:  python -m torch.distributed.run --nproc-per-node 8 examples/synthetic_multi_gpu.py
(every rank runs the same seeded GA; each generation's individuals are sharded and the fitness all-gathered)."""
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

if __name__ == '__main__':
    import gentun_b200 as gentun
    from gentun_b200.sharding import init_distributed
    from tests.synth import make_dataset

    generations = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    size = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    rows = int(sys.argv[3]) if len(sys.argv) > 3 else 1_000_000
    rank, world = (0, 1)
    if "RANK" in os.environ:
        rank, world = init_distributed()
    x, y = make_dataset(rows, 50, seed=1003)
    random.seed(20240601)
    if rank != 0:
        sys.stdout = open(os.devnull, "w")
    pop = gentun.Population(gentun.XgboostIndividual, x, y, size=size, maximize=False,
                            additional_parameters={'nfold': 5, 'num_boost_round': 100})
    ga = gentun.GeneticAlgorithm(pop)
    t0 = time.time()
    for g in range(generations):
        t1 = time.time()
        ga.evolve_population()
        ga.generation += 1
        print("generation %d: %.2f s on %d GPU(s)" % (g + 1, time.time() - t1, world))
    print("total %.2f s" % (time.time() - t0))
