"""configs[4] (Rotamers on 12 side chains, 32,768 individuals) step time on one GPU, with a checksum
of the scores so that two builds can be compared on the same box (tools/ab_compare.sh)."""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import bench

batch, tors, xf, rot = bench.build_c5(0)
eng = batch.engines[0][0]
dev = eng.torch_device
flush = torch.empty(192 << 20, dtype=torch.uint8, device=dev)
for pop in (4096, 32768):
    t, x, r = (torch.as_tensor(v[:pop]).to(dev) for v in (tors, xf, rot))
    sc = torch.empty((pop, 2), dtype=torch.float64, device=dev)
    cn = torch.empty((pop, 2), dtype=torch.int32, device=dev)
    fn = lambda: eng.evaluate(t, x, r, scores_out=sc, counts_out=cn)
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    for _ in range(3):
        g.replay()
    ts = []
    for k in range(10):
        flush.fill_(k & 255)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    digest = hashlib.sha1(sc.cpu().numpy().tobytes() + cn.cpu().numpy().tobytes()).hexdigest()[:12]
    print("configs4 pop %5d graph %.4f ms (min %.4f)  scores %s" % (pop, float(np.median(ts)), min(ts), digest))
