#!/usr/bin/env python
"""Small driver for profiling one configuration under ncu:
    python tools/prof_case.py dense|culled|tiled|rotamers [pop]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine

mode = sys.argv[1] if len(sys.argv) > 1 else "culled"
pop = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
if mode == "rotamers":          # BASELINE configs[4] through bench.py's builder
    import bench
    batch, tors, xf, rot = bench.build_c5(0)
    eng = batch.engines[0][0]
    eng.set_profiling(True)
    n = min(pop, len(tors))
    t, x, r = (torch.as_tensor(v[:n]).cuda() for v in (tors, xf, rot))
    for _ in range(4):
        eng.evaluate(t, x, r)
    torch.cuda.synchronize()
    print(mode, n, eng.kernel_times(), eng.last_pair_tests())
    sys.exit(0)
params = M.ContactParams(cull=(mode != "dense"), tiled=(mode == "tiled"))
lig = synth.make_ligand(40, 10, seed=1)
rec = synth.make_receptor(20000, seed=2, params=params)
tors, xf = synth.make_alleles(pop, 10, seed=3)
eng = Engine(lig, rec, params)
eng.set_profiling(True)
t, x = torch.as_tensor(tors).cuda(), torch.as_tensor(xf).cuda()
for _ in range(4):
    eng.evaluate(t, x)
torch.cuda.synchronize()
print(mode, pop, eng.kernel_times(), eng.last_pair_tests())
