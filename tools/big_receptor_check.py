import sys, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine
for R in (120000, 400000):
    params = M.ContactParams()
    lig = synth.make_ligand(40, 10, seed=1)
    t0 = time.time(); rec = synth.make_receptor(R, seed=2, params=params); t1 = time.time()
    eng = Engine(lig, rec, params); t2 = time.time()
    tiled = Engine(lig, rec, M.ContactParams(tiled=True))
    tors, xf = synth.make_alleles(4096, 10, seed=3, radius=20.0)
    eng.set_profiling(True)
    for _ in range(3): s, c, _ = eng.evaluate(tors, xf)
    torch.cuda.synchronize(); kt = eng.kernel_times()
    s2, c2, _ = tiled.evaluate(tors, xf)
    print("R %d: build %.1fs, context %.1fs, fused %.3f ms for 4096, identical to tiled: %s, nonzero %d" % (
        R, t1 - t0, t2 - t1, kt["sweep"], bool(torch.equal(s, s2) and torch.equal(c, c2)), int((s[:, 1] != 0).sum())))
    eng.close(); tiled.close()
