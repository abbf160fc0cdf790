#!/usr/bin/env python
"""Kernel time against population size on configs[2]'s model: where the partly filled last wave
of warp-per-individual work starts to cost (8 GPUs leave 8,192 individuals per GPU).
    python tools/tail_curve.py [pop ...]
Prints, per population: express and Contacts kernel time with the L2 flushed before each launch
and with a warm L2, and the per-individual cost relative to the 65,536 run."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine

pops = [int(v) for v in sys.argv[1:]] or [2368, 4736, 7104, 8192, 9472, 11840, 16384, 32768, 65536]
params = M.ContactParams()
lig = synth.make_ligand(40, 10, seed=1)
rec = synth.make_receptor(20000, seed=2, params=params)
tors, xf = synth.make_alleles(65536, 10, seed=3)
eng = Engine(lig, rec, params)
eng.reserve(65536)
eng.set_profiling(True)
t, x = torch.as_tensor(tors).cuda(), torch.as_tensor(xf).cuda()
flush = torch.empty(192 * 1024 * 1024, dtype=torch.uint8, device="cuda")
rows = {}
for pop in pops:
    res = {}
    for mode in ("flushed", "warm"):
        ks, es = [], []
        for k in range(12):
            if mode == "flushed":
                flush.fill_(k & 0xff)
            eng.evaluate(t[:pop], x[:pop])
            kt = eng.kernel_times()
            if k >= 2:
                ks.append(kt["sweep"]); es.append(kt["express"])
        res[mode] = (float(np.median(es)), float(np.median(ks)))
    rows[pop] = res
ref = rows[max(pops)]["flushed"][1] / max(pops)
for pop in pops:
    r = rows[pop]
    print("pop %6d  flushed: express %.4f contacts %.4f ms | warm: express %.4f contacts %.4f ms | "
          "per-individual cost x%.3f of the largest run" % (pop, r["flushed"][0], r["flushed"][1], r["warm"][0],
                                                           r["warm"][1], r["flushed"][1] / pop / ref))
