#!/usr/bin/env python
"""Share of set bits in the configs[3] similarity bitmask (what a CSR partner list would carry)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
lig = synth.make_ligand(12, 2, seed=1)
conf = torch.as_tensor(synth.make_conformers(n, 60, seed=4).astype(np.float32)).cuda()
eng = Engine(lig, M.empty_receptor(), M.ContactParams(only_internal=True))
_, bits = eng.rmsd_matrix(conf, threshold=2.5, values=False)
ptr, idx = eng.similar_lists(bits, n)
print("N %d: %d partner entries (%.4f %% of N^2), CSR bytes %d vs bitmask bytes %d" %
      (n, len(idx), 100.0 * len(idx) / n / n, idx.nbytes + ptr.nbytes, n * ((n + 31) // 32) * 4))
