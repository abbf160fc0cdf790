#!/usr/bin/env python
"""Wall time of the plugin-path GA (per-individual objects, `gaudi_b200.algorithms.launch`) on the
5er1 test complex: BASELINE.json configs[0] (population 100 x 10 generations) and ten times that."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_gpu_plugins as T
from gaudi_b200 import algorithms
algorithms.launch(T.job(100, 1), seed=1)          # library load, context set-up
for pop, gens in ((100, 10), (1000, 10), (5000, 5)):
    t0 = time.time()
    p, log, elite = algorithms.launch(T.job(pop, gens), seed=2)
    print("population %5d x %2d generations: %.2f s   best clashes %.3f" % (
        pop, gens, time.time() - t0, min(i.fitness.values[0] for i in p)))
