#!/usr/bin/env python
"""Does splitting a small shard into chunks on two streams (express of chunk k+1 under Contacts of
chunk k, Contacts kernels draining into each other) shorten the step?  Population 8,192 (what one of
8 GPUs gets of configs[2]) and 16,384."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from gaudi_b200 import model as M, synth
from gaudi_b200.engine import Engine

params = M.ContactParams()
lig = synth.make_ligand(40, 10, seed=1)
rec = synth.make_receptor(20000, seed=2, params=params)
tors, xf = synth.make_alleles(65536, 10, seed=3)
eng = Engine(lig, rec, params)
eng.reserve(65536)
t, x = torch.as_tensor(tors).cuda(), torch.as_tensor(xf).cuda()
flush = torch.empty(192 << 20, dtype=torch.uint8, device="cuda")
s2 = torch.cuda.Stream()

def timed(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    out = []
    for k in range(reps):
        flush.fill_(k & 255)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        out.append(a.elapsed_time(b))
    return float(np.median(out))

for pop in (8192, 16384):
    xyz = torch.empty((pop, 40, 3), dtype=torch.float32, device="cuda")
    sc = torch.empty((pop, 2), dtype=torch.float64, device="cuda")
    cn = torch.empty((pop, 2), dtype=torch.int32, device="cuda")
    def whole():
        eng.evaluate(t[:pop], x[:pop], scores_out=sc, counts_out=cn, xyz_out=xyz)
    def graph_of(fn):
        fn(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g): fn()
        return g
    def chunks(n):
        b = [pop * k // n for k in range(n + 1)]
        def fn():
            main = torch.cuda.current_stream()
            ev = []
            # express of every chunk on the side stream, back to back; Contacts on the main stream
            s2.wait_stream(main)
            with torch.cuda.stream(s2):
                for k in range(n):
                    eng.express(t[b[k]:b[k+1]], x[b[k]:b[k+1]], out=xyz[b[k]:b[k+1]])
                    e = torch.cuda.Event(); e.record(); ev.append(e)
            for k in range(n):
                main.wait_event(ev[k])
                import ctypes as C
                from gaudi_b200 import _lib
                _lib.check(eng._L.gmm_contacts(eng._h, b[k+1]-b[k], C.c_void_p(xyz[b[k]:b[k+1]].data_ptr()),
                           C.c_void_p(sc[b[k]:b[k+1]].data_ptr()), C.c_void_p(cn[b[k]:b[k+1]].data_ptr()), eng._stream()))
        return fn
    g1 = graph_of(whole)
    print("pop %5d  whole eager %.4f ms  graph %.4f ms" % (pop, timed(whole), timed(g1.replay)))
    ref = sc.clone()
    for n in (2, 3, 4):
        f = chunks(n)
        g = graph_of(f)
        ms_e, ms_g = timed(f), timed(g.replay)
        torch.cuda.synchronize()
        print("          %d chunks eager %.4f ms  graph %.4f ms  same scores %s" % (n, ms_e, ms_g, bool(torch.equal(sc, ref))))
