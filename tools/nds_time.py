#!/usr/bin/env python
"""Device non-dominated sort timing: N rows x M objectives (gmm_nondominated_fronts, incl. the host
lexsort and the copies)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine
eng = Engine(synth.make_ligand(12, 2, seed=1), M.empty_receptor(), M.ContactParams(only_internal=True))
rng = np.random.default_rng(0)
for n, m in ((4000, 3), (16000, 3), (131072, 2), (131072, 3), (131072, 4)):
    w = np.round(rng.normal(size=(n, m)) * 50, 3)
    eng.nondominated_fronts(w[:1000])
    t0 = time.perf_counter()
    f = eng.nondominated_fronts(w)
    print("N %6d  M %d  fronts %4d  %.1f ms" % (n, m, f.max() + 1, (time.perf_counter() - t0) * 1e3))
