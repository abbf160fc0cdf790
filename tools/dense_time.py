import os, sys
sys.path.insert(0, '/root/repo')
import torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine
params = M.ContactParams(cull=False)
lig = synth.make_ligand(40, 10, seed=1); rec = synth.make_receptor(20000, seed=2, params=params)
tors, xf = synth.make_alleles(16384, 10, seed=3)
eng = Engine(lig, rec, params); eng.set_profiling(True)
t, x = torch.as_tensor(tors).cuda(), torch.as_tensor(xf).cuda()
for _ in range(4): eng.evaluate(t, x)
torch.cuda.synchronize(); print("dense 16384", eng.kernel_times())
