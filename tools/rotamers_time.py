#!/usr/bin/env python
"""BASELINE.json configs[4] on one GPU (12 flexible side chains, population 32,768): bench.py's side measurement alone."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


r = bench.rotamers_measurement(torch, torch.device("cuda", 0), timed)
print({k: r[k] for k in ("value", "ms_per_step", "express_ms", "contacts_ms", "mobile_atoms")})
