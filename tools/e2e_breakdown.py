#!/usr/bin/env python
"""Where the end-to-end (host buffers in, host buffers out) time of one evaluate goes:
H2D of the alleles alone, the kernels alone, D2H of the scores alone, and gmm_evaluate_host."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine

pop = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
params = M.ContactParams()
lig = synth.make_ligand(40, 10, seed=1)
rec = synth.make_receptor(20000, seed=2, params=params)
eng = Engine(lig, rec, params)
eng.reserve(pop)
if os.environ.get('PROF'):
    eng.set_profiling(True)
tors, xf = synth.make_alleles(pop, 10, seed=3)
th, xh = torch.as_tensor(tors).pin_memory(), torch.as_tensor(xf).pin_memory()
td, xd = th.cuda(), xh.cuda()
s_pin = torch.empty((pop, 2), dtype=torch.float64).pin_memory()
c_pin = torch.empty((pop, 2), dtype=torch.int32).pin_memory()
sd = torch.empty((pop, 2), dtype=torch.float64, device="cuda")
cd = torch.empty((pop, 2), dtype=torch.int32, device="cuda")


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
        torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


def h2d():
    td.copy_(th, non_blocking=True); xd.copy_(xh, non_blocking=True)


def d2h():
    s_pin.copy_(sd, non_blocking=True); c_pin.copy_(cd, non_blocking=True)


print("pop %d: H2D %.3f ms (%.1f GB/s)  kernels %.3f ms  D2H %.3f ms  evaluate_host (pinned) %.3f ms  evaluate_host (pageable) %.3f ms" % (
    pop, timed(h2d), (th.numel() + xh.numel()) * 8 / timed(h2d) / 1e6,
    timed(lambda: eng.evaluate(td, xd, scores_out=sd, counts_out=cd)), timed(d2h),
    timed(lambda: eng.evaluate_host(th, xh, scores_out=s_pin, counts_out=c_pin)),
    timed(lambda: eng.evaluate_host(tors, xf))))
