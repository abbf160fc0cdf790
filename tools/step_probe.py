import os, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine
params = M.ContactParams(); lig = synth.make_ligand(40, 10, seed=1); rec = synth.make_receptor(20000, seed=2, params=params)
tors, xf = synth.make_alleles(65536, 10, seed=3)
eng = Engine(lig, rec, params); eng.reserve(65536)
t, x = torch.as_tensor(tors).cuda(), torch.as_tensor(xf).cuda()
flush = torch.empty(192 << 20, dtype=torch.uint8, device="cuda")
for pop in (8192, 16384, 65536):
    sc = torch.empty((pop, 2), dtype=torch.float64, device="cuda"); cn = torch.empty((pop, 2), dtype=torch.int32, device="cuda")
    fn = lambda: eng.evaluate(t[:pop], x[:pop], scores_out=sc, counts_out=cn)
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): fn()
    for name, f in (("eager", fn), ("graph", g.replay)):
        for _ in range(3): f()
        ts = []
        for k in range(20):
            flush.fill_(k & 255)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); f(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
        print("pop %5d %s %.4f ms (min %.4f)" % (pop, name, float(np.median(ts)), min(ts)))
    ref = eng.evaluate(t[:pop], x[:pop])[0]
    g.replay(); torch.cuda.synchronize()
    print("   same scores:", bool(torch.equal(ref, sc)))
