"""Coordination objective throughput on one GPU: 4c3w (the reference's test complex), the vanadium
complex moved by Search alleles near the crystal pose, octahedral template.  Prints the time of
gmm_coordination (sphere search + CPD kernels) per population and the CPU oracle's rate on a sample."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import oracle_adapter as OA
import test_gpu_coordination as T

mol, site, g, eng, mob = T.complex_engine()
for pop in (4096, 65536):
    xf = T.poses(mob, pop, seed=7)
    xyz = eng.express(np.zeros((pop, 0)), xf)
    eng.coordination(xyz); torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); out = eng.coordination(xyz); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    ms = float(np.median(ts))
    print("coordination pop %6d: %.3f ms  (%.1f M individuals/s)" % (pop, ms, pop / ms / 1e3))
x64 = xyz[:300].cpu().numpy().astype(np.float64) + eng.centre
t0 = time.perf_counter()
for p in range(300):
    OA.coordination(site, OA.coordination_world(site, x64[p]))
dt = time.perf_counter() - t0
print("oracle (numpy, one core): %.1f individuals/s" % (300 / dt))
