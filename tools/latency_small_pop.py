#!/usr/bin/env python
"""Evaluate latency for the population sizes GaudiMM jobs usually run (docs/faq.rst:48-50)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine
params = M.ContactParams()
lig = synth.make_ligand(40, 10, seed=1)
rec = synth.make_receptor(20000, seed=2, params=params)
eng = Engine(lig, rec, params)
eng.set_profiling(True)
for pop in [int(v) for v in (sys.argv[1:] or (1, 10, 100, 300, 1000, 1183, 1184, 4096))]:
    tors, xf = synth.make_alleles(pop, 10, seed=3)
    t, x = torch.as_tensor(tors).cuda(), torch.as_tensor(xf).cuda()
    for _ in range(5):
        eng.evaluate(t, x)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50):
        eng.evaluate(t, x)
    b.record(); torch.cuda.synchronize()
    dev_ms = a.elapsed_time(b) / 50
    t0 = time.perf_counter()
    for _ in range(50):
        eng.evaluate_host(tors, xf)
    host_ms = (time.perf_counter() - t0) / 50 * 1e3
    print("pop %5d  device %.3f ms (%.0f ind/s)  host-to-host %.3f ms  kernels %s" % (
        pop, dev_ms, pop / dev_ms * 1e3, host_ms, {k: round(v, 4) for k, v in eng.kernel_times().items()}))
