"""Where the time of parallel.ShardedSimilarity.step goes on one GPU (a synchronise after every part;
131,072 conformers by default):  python tools/sim_step_probe.py [N]"""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from gaudi_b200 import _lib, model as M, parallel, synth
from gaudi_b200.engine import Engine

N = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
A = 60
lig = synth.make_ligand(12, 2, seed=1)
eng = Engine(lig, M.empty_receptor(), M.ContactParams(only_internal=True), device=0)
conf = synth.make_conformers(16384, A, seed=4).astype(np.float32)
if N > 16384:
    rng = np.random.default_rng(5)
    conf = np.concatenate([conf + rng.normal(scale=0.5, size=(1, 1, 3)).astype(np.float32) * k for k in range(N // 16384)])
coords = torch.as_tensor(conf[:N]).cuda()
sim = parallel.ShardedSimilarity(eng, [N], A, threshold=2.5)
for _ in range(3):
    ptr, idx = sim.step(coords)
print("N %d: %d partner entries, cap %d" % (N, len(idx), sim.cap))
t = []
for _ in range(5):
    torch.cuda.synchronize(); t0 = time.perf_counter(); sim.step(coords); t.append((time.perf_counter() - t0) * 1e3)
print("step %.3f ms (min %.3f)" % (float(np.median(t)), min(t)))
# the parts, with a synchronise after each
L, h, stream = eng._L, eng._h, eng._stream()
p = lambda x: C.c_void_p(x.data_ptr())
def timed(name, fn, reps=5):
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter(); out = fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    print("  %-34s %.3f ms" % (name, float(np.median(ts))))
    return out
timed("bitmask (centre + tcgen05 kernel)", lambda: _lib.check(L.gmm_rmsd_matrix(h, N, A, p(coords), 0, N, None, p(sim.bits), 2.5, stream)))
cd = sim.packed[:N]
timed("count", lambda: _lib.check(L.gmm_similar_lists(h, N, 0, N, p(sim.bits), p(cd), None, None, 0, stream)))
timed("cumsum", lambda: torch.cumsum(cd, 0, out=sim.ptr[1:]))
timed("fill", lambda: _lib.check(L.gmm_similar_lists(h, N, 0, N, p(sim.bits), None, p(sim.ptr), C.c_void_p(sim.packed.data_ptr() + 4 * sim.m), sim.cap, stream)))
timed("D2H of the counts", lambda: sim.host_counts.copy_(sim.packed.view(1, sim.m + sim.cap)[:, :sim.m], non_blocking=True))
tot = int(sim.host_counts.numpy()[0, :N].sum())
timed("D2H of the entries (%.1f MB)" % (tot * 4 / 1e6), lambda: sim.host_idx[:tot].copy_(sim.packed[sim.m:sim.m + tot], non_blocking=True))
def host_part():
    ptr = np.zeros(N + 1, np.int64)
    np.cumsum(sim.host_counts.numpy()[0, :N], out=ptr[1:])
    return ptr
timed("host: row pointers (numpy cumsum)", host_part)
