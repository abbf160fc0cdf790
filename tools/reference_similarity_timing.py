#!/usr/bin/env python
"""BUILD CONTAINER ONLY: times the reference's own gaudi/similarity.py (_rmsd_squared, called pair by
pair as gaudi/algorithms.py:141-155 does) beside the oracle's restatement of the same three lines, on
the 1,024-conformer subsample bench.py uses.  /root/reference does not exist on the GPU box, so this
cannot be part of bench.py; its output is committed as profiles/reference_similarity_timing_r02.txt."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, "/root/reference")
import numpy as np
from gaudi.similarity import _rmsd_squared as ref_fn          # the reference's own function
import gaudi_oracle as O
from gaudi_b200 import synth
sub = synth.make_conformers(16384, 60, seed=4).astype(np.float32).astype(np.float64)[:1024]
for name, fn in (("reference gaudi.similarity._rmsd_squared", ref_fn), ("oracle.rmsd_squared (restatement)", O.rmsd_squared)):
    t0, n, acc = time.perf_counter(), 0, 0
    for i in range(len(sub)):
        a = sub[i]
        for j in range(i + 1, len(sub)):
            acc += not (fn(a, sub[j]) > 6.25)
        n += len(sub) - 1 - i
    dt = time.perf_counter() - t0
    print("%-45s %8d pairs in %.2f s = %.0f pairs/s (%d similar), %d core(s) of %d in this container"
          % (name, n, dt, n / dt, acc, 1, os.cpu_count()))
