#!/usr/bin/env python
"""Similarity-bitmask timing of BASELINE.json configs[3] (16,384 conformers x 60 atoms): tensor-core
pre-filter path against the FP32 direct-difference kernel, and a bit-for-bit comparison."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
lig = synth.make_ligand(12, 2, seed=1)
conf = torch.as_tensor(synth.make_conformers(n, 60, seed=4).astype(np.float32)).cuda()
out = {}
for name, simt in (("mma", False), ("simt", True)):
    eng = Engine(lig, M.empty_receptor(), M.ContactParams(only_internal=True, rmsd_simt=simt))
    for _ in range(3):
        _, bits = eng.rmsd_matrix(conf, threshold=2.5, values=False)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        _, bits = eng.rmsd_matrix(conf, threshold=2.5, values=False)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    out[name] = bits.clone()
    print("%-5s %.3f ms  %.1f G unique pairs/s" % (name, ms, n * (n - 1) / 2 / ms / 1e6))
print("identical bits:", bool(torch.equal(out["mma"], out["simt"])))
