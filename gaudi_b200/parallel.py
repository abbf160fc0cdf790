"""Multi-GPU plumbing: one process per GPU (``torchrun``), population sharded by individual,
one ``all_gather`` of the fitness block per evaluation.  The data path
has no collective: individuals are independent (gaudi/algorithms.py:99-102 is a pure map).
RMSD: coordinates are all-gathered, each rank computes a block of rows.

Works with the ``nccl`` backend on GPUs and with ``gloo`` on CPU tensors (used by the
world-size-2 tests, where the per-shard compute is supplied by the test)."""
from __future__ import annotations

import torch
import torch.distributed as dist


def _all_gather_into(out, inp, group=None):
    """``all_gather_into_tensor``; device tensors under a gloo group (tests with two ranks on one
    GPU, where NCCL cannot run) are staged through the host."""
    if inp.is_cuda and dist.get_backend(group) == "gloo":
        staged = torch.empty(out.shape, dtype=out.dtype)
        dist.all_gather_into_tensor(staged, inp.cpu(), group=group)
        out.copy_(staged)
    else:
        dist.all_gather_into_tensor(out, inp, group=group)


def shard_bounds(n, world):
    """Contiguous, balanced blocks: rank r owns [b[r], b[r+1])."""
    return [(n * r) // world for r in range(world + 1)]


def max_shard(n, world):
    b = shard_bounds(n, world)
    return max(b[r + 1] - b[r] for r in range(world))


def sharded_evaluate(evaluate_shard, alleles, n_out=2, device=None, group=None):
    """Evaluate a population split over the ranks of ``group``.

    ``alleles``: tuple of arrays with the population on axis 0 (every rank holds all of them —
    they are tiny — or at least its own slice); ``evaluate_shard(*slices) -> (p, n_out)`` tensor
    or array for this rank's slice.  Returns the full ``(P, n_out)`` float64 tensor on every
    rank (rank 0 is the one that runs NSGA-II selection)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    P = int(alleles[0].shape[0])
    b = shard_bounds(P, world)
    lo, hi = b[rank], b[rank + 1]
    part = evaluate_shard(*[a[lo:hi] if a is not None else None for a in alleles])
    part = torch.as_tensor(part, dtype=torch.float64)
    if device is not None:
        part = part.to(device)
    if world == 1:
        return part
    m = max_shard(P, world)
    padded = torch.zeros((m, n_out), dtype=torch.float64, device=part.device)
    padded[:hi - lo] = part
    gathered = torch.empty((world * m, n_out), dtype=torch.float64, device=part.device)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    out = torch.cat([gathered[r * m:r * m + (b[r + 1] - b[r])] for r in range(world)])
    return out


def sharded_rmsd_bits(rmsd_rows, coords, group=None, counts=None):
    """Row-block sharded similarity bitmask.  ``coords``: this rank's (n_r, A, 3) float32 block
    of conformers; ``rmsd_rows(all_coords, row0, row1)`` returns the (row1-row0, ceil(N/32)) int32
    words for those rows.  ``counts``: conformers per rank when the caller knows them (blocks that
    follow ``shard_bounds``); without it the ranks exchange their block sizes first, which costs a
    host round trip per rank.  Returns the full (N, words) tensor on every rank."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1:
        return rmsd_rows(coords, 0, coords.shape[0])
    if counts is None:
        sizes = torch.zeros(world, dtype=torch.int64, device=coords.device)
        dist.all_gather_into_tensor(sizes, torch.tensor([coords.shape[0]], dtype=torch.int64, device=coords.device),
                                    group=group)
        counts = sizes.tolist()                     # one device-to-host copy
    counts = [int(c) for c in counts]
    m = max(counts)
    even = min(counts) == m                         # equal blocks: gather in place, no padding, no re-packing
    if even:
        padded = coords.contiguous()
    else:
        padded = torch.zeros((m,) + tuple(coords.shape[1:]), dtype=coords.dtype, device=coords.device)
        padded[:coords.shape[0]] = coords
    allc = torch.empty((world * m,) + tuple(coords.shape[1:]), dtype=coords.dtype, device=coords.device)
    dist.all_gather_into_tensor(allc, padded, group=group)
    if not even:
        allc = torch.cat([allc[r * m:r * m + counts[r]] for r in range(world)])
    n = allc.shape[0]
    row0 = sum(counts[:rank])
    mine = rmsd_rows(allc, row0, row0 + counts[rank])
    words = (n + 31) // 32
    if even:
        pad = mine
    else:
        pad = torch.zeros((m, words), dtype=mine.dtype, device=mine.device)
        pad[:counts[rank]] = mine
    allb = torch.empty((world * m, words), dtype=mine.dtype, device=mine.device)
    dist.all_gather_into_tensor(allb, pad, group=group)
    if even:
        return allb
    return torch.cat([allb[r * m:r * m + counts[r]] for r in range(world)])


def sharded_similar_lists(engine, coords, counts, threshold=2.5, group=None):
    """Row-block sharded similarity (gaudi/similarity.py:36-75 over all pairs, the input of the
    niching pass gaudi/algorithms.py:139-155) where only the PARTNER LISTS travel: ``coords`` is this
    rank's ``(n_r, A, 3)`` float32 block of conformers (``counts[r]`` conformers on rank r); the
    conformers are all-gathered (the one real exchange of this stage: every rank needs every
    column), each rank computes its rows of the bitmask and compacts them into partner lists on its
    own GPU, and the lists — a few hundred KB where the bitmask is tens of MB — are gathered.
    Returns ``(ptr int64 (N + 1,), idx int32)`` as device tensors on every rank; the bitmask rows stay
    where they were computed."""
    world, rank = _world(group)
    counts = [int(c) for c in counts]
    n = sum(counts)
    row0 = sum(counts[:rank])
    if world == 1:
        _, bits = engine.rmsd_matrix(coords, threshold=threshold, values=False)
        return engine.similar_lists(bits, n, to_host=False)
    m = max(counts)
    even = min(counts) == m
    if even:
        padded = coords.contiguous()
    else:
        padded = torch.zeros((m,) + tuple(coords.shape[1:]), dtype=coords.dtype, device=coords.device)
        padded[:coords.shape[0]] = coords
    allc = torch.empty((world * m,) + tuple(coords.shape[1:]), dtype=coords.dtype, device=coords.device)
    _all_gather_into(allc, padded, group)
    if not even:
        allc = torch.cat([allc[r * m:r * m + counts[r]] for r in range(world)])
    _, bits = engine.rmsd_matrix(allc, row0=row0, row1=row0 + counts[rank], threshold=threshold, values=False)
    ptr, idx = engine.similar_lists(bits, n, row0=row0, to_host=False)
    # per-row partner counts of every rank (fixed size), then the list entries padded to the longest
    per_row = torch.zeros(m, dtype=torch.int64, device=coords.device)
    per_row[:counts[rank]] = ptr[1:] - ptr[:-1]
    all_rows = torch.empty(world * m, dtype=torch.int64, device=coords.device)
    _all_gather_into(all_rows, per_row, group)
    all_rows = all_rows.view(world, m)
    totals = all_rows.sum(dim=1).tolist()                 # one device-to-host copy
    cap = max(max(totals), 1)
    mine = torch.zeros(cap, dtype=torch.int32, device=coords.device)
    mine[:idx.shape[0]] = idx
    all_idx = torch.empty(world * cap, dtype=torch.int32, device=coords.device)
    _all_gather_into(all_idx, mine, group)
    rows = torch.cat([all_rows[r, :counts[r]] for r in range(world)])
    full_ptr = torch.zeros(n + 1, dtype=torch.int64, device=coords.device)
    torch.cumsum(rows, 0, out=full_ptr[1:])
    full_idx = torch.cat([all_idx[r * cap:r * cap + int(totals[r])] for r in range(world)])
    return full_ptr, full_idx


class ShardedSimilarity(object):
    """``sharded_similar_lists`` for a generational loop: every buffer is allocated once, a step is
    two collectives (conformers in, packed partner lists out), nine kernel launches and two host
    synchronisations (the per-row counts first, then exactly the entries they announce, copied
    straight to their final place in one page-locked array), and the partner lists arrive as HOST
    arrays (the greedy niching pass that consumes them, gaudi/algorithms.py:139-155, runs on the
    host) — on every rank, or with ``root_only`` on rank 0 alone.  The arrays returned are views of
    buffers the next ``step`` reuses.

    ``counts[r]`` conformers of ``atoms`` atoms live on rank r.  ``cap``: room for partner entries per
    rank (default 8 per row: the whole packed buffer crosses PCIe every step, so it is kept small); a
    step that needs more doubles it, runs again, and the larger buffer stays."""

    def __init__(self, engine, counts, atoms, threshold=2.5, group=None, cap=None):
        import numpy as np
        self.engine, self.group, self.threshold = engine, group, float(threshold)
        self.world, self.rank = _world(group)
        self.counts = [int(c) for c in counts]
        self.n, self.m, self.atoms = sum(self.counts), max(self.counts), int(atoms)
        self.row0 = sum(self.counts[:self.rank])
        self.rows = self.counts[self.rank]
        dev = engine.torch_device
        self.even = min(self.counts) == self.m
        self.allc = torch.empty((self.world * self.m, self.atoms, 3), dtype=torch.float32, device=dev)
        self.pad = None if self.even else torch.zeros((self.m, self.atoms, 3), dtype=torch.float32, device=dev)
        self.bits = torch.empty((self.rows, (self.n + 31) // 32), dtype=torch.int32, device=dev)
        self.ptr = torch.zeros(self.rows + 1, dtype=torch.int64, device=dev)
        self._np = np
        self._alloc(int(cap) if cap else 8 * self.m)
        engine.reserve_rmsd(self.n, self.atoms)

    def _alloc(self, cap):
        dev = self.engine.torch_device
        self.cap = cap
        self.packed = torch.zeros(self.m + cap, dtype=torch.int32, device=dev)      # [per-row counts | entries]
        self.gathered = torch.empty(self.world * (self.m + cap), dtype=torch.int32, device=dev)
        self.host_counts = torch.empty((self.world, self.m), dtype=torch.int32).pin_memory()
        self.host_idx = torch.empty(self.world * cap, dtype=torch.int32).pin_memory()

    def step(self, coords, root_only=False):
        """``coords``: this rank's ``(counts[rank], atoms, 3)`` float32 device block.  Returns host
        ``(ptr int64 (N + 1,), idx int32)`` for all N conformers; with ``root_only`` the ranks other
        than 0 return ``(None, None)`` and skip the copy of the entries."""
        import ctypes as C
        from . import _lib
        eng, np = self.engine, self._np
        L, h, stream = eng._L, eng._h, eng._stream()
        if self.world > 1:
            src = coords
            if not self.even:
                self.pad[:self.rows] = coords
                src = self.pad
            _all_gather_into(self.allc, src.contiguous(), self.group)
            allc = self.allc
            if not self.even:
                allc = torch.cat([self.allc[r * self.m:r * self.m + self.counts[r]] for r in range(self.world)])
        else:
            allc = coords.contiguous()
        p = lambda t: C.c_void_p(t.data_ptr())
        _lib.check(L.gmm_rmsd_matrix(h, self.n, self.atoms, p(allc), self.row0, self.row0 + self.rows, None, p(self.bits),
                                     self.threshold, stream))
        counts_d = self.packed[:self.rows]
        _lib.check(L.gmm_similar_lists(h, self.n, self.row0, self.row0 + self.rows, p(self.bits), p(counts_d), None, None, 0, stream))
        torch.cumsum(counts_d, 0, out=self.ptr[1:])
        # entries beyond the capacity are dropped by the bound the fill pass is given; the host sees
        # the true totals in the counts and retries with more room
        _lib.check(L.gmm_similar_lists(h, self.n, self.row0, self.row0 + self.rows, p(self.bits), None, p(self.ptr),
                                       C.c_void_p(self.packed.data_ptr() + 4 * self.m), self.cap, stream))
        if self.world > 1:
            _all_gather_into(self.gathered, self.packed, self.group)
        rows = (self.gathered if self.world > 1 else self.packed).view(self.world, self.m + self.cap)
        self.host_counts.copy_(rows[:, :self.m], non_blocking=True)
        stream_t = torch.cuda.current_stream(eng.torch_device)
        stream_t.synchronize()
        counts_h = self.host_counts.numpy()
        per_row = [counts_h[r, :self.counts[r]] for r in range(self.world)]
        totals = [int(c.sum()) for c in per_row]
        if max(totals) > self.cap:                            # every rank sees the same counts and retries
            self._alloc(2 * max(totals))
            return self.step(coords, root_only)
        shrink = self.cap > 4 * max(max(totals), self.m)      # far too roomy: the all-gather pays for it
        ptr = idx = None
        if not (root_only and self.rank != 0):
            at = 0
            for r in range(self.world):
                if totals[r]:
                    self.host_idx[at:at + totals[r]].copy_(rows[r, self.m:self.m + totals[r]], non_blocking=True)
                at += totals[r]
            ptr = np.zeros(self.n + 1, np.int64)
            np.cumsum(np.concatenate(per_row), out=ptr[1:])
            stream_t.synchronize()
            idx = self.host_idx.numpy()[:at]
        if shrink:
            if idx is not None:
                idx = idx.copy()                              # (the results leave the old buffer first)
            self._alloc(2 * max(max(totals), self.m))
        return ptr, idx


# ---------------------------------------------------------------------------------------------
# Fitness exchange through peer memory (gmm_exchange_*): the all-gather without a collective
# ---------------------------------------------------------------------------------------------
def _world(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group), dist.get_rank(group)
    return 1, 0


def tensor_from_pointer(ptr, shape, dtype, device):
    """Zero-copy torch view of device memory owned by the library (``__cuda_array_interface__``)."""
    typestr = {torch.float64: "<f8", torch.int32: "<i4", torch.float32: "<f4"}[dtype]
    holder = type("DevicePointer", (), {})()
    holder.__cuda_array_interface__ = {"shape": tuple(int(s) for s in shape), "typestr": typestr,
                                       "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(holder, device=device)


class FitnessExchange(object):
    """Every rank's ``[n, 2]`` score rows (and pair counts) land in every rank's copy of the
    ``[max_rows, 2]`` fitness table: one small kernel per step that stores through CUDA-IPC peer
    mappings, publishes an epoch flag and waits for the other ranks' flags (csrc/exchange.cu).  The
    process group is used once, at construction, to pass the IPC handles around; a step contains no
    collective call and can be replayed from a CUDA graph."""

    def __init__(self, engine, max_rows, group=None):
        import ctypes as C
        from . import _lib
        self.engine, self.group, self.max_rows = engine, group, int(max_rows)
        self.world, self.rank = _world(group)
        self._L, self._C = _lib.lib(), C
        handle = (C.c_ubyte * 64)()
        _lib.check(self._L.gmm_exchange_create(engine._h, self.world, self.rank, self.max_rows, handle))
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle), group=group)
            blob = (C.c_ubyte * (64 * self.world)).from_buffer_copy(b"".join(handles))
            _lib.check(self._L.gmm_exchange_connect(engine._h, blob))
            dist.barrier(group=group)           # every rank has mapped every block before anyone publishes

    def _views(self, ps, pc):
        dev = self.engine.torch_device
        return (tensor_from_pointer(ps.value, (self.max_rows, 2), torch.float64, dev),
                tensor_from_pointer(pc.value, (self.max_rows, 2), torch.int32, dev))

    def publish(self, scores, counts, row0):
        """Enqueue the exchange of this rank's rows on torch's current stream; returns the
        ``(scores [max_rows, 2], counts [max_rows, 2])`` views that hold every rank's rows once the
        kernel has run (valid until this rank's publish after next)."""
        from . import _lib
        C = self._C
        ps, pc = C.c_void_p(), C.c_void_p()
        _lib.check(self._L.gmm_exchange_publish(
            self.engine._h, C.c_void_p(scores.data_ptr()), C.c_void_p(counts.data_ptr()) if counts is not None else None,
            int(row0), int(scores.shape[0]), C.byref(ps), C.byref(pc), self.engine._stream()))
        return self._views(ps, pc)

    def evaluate_publish(self, torsions, xforms, rot_alleles, scores, counts, row0):
        """``Engine.evaluate`` into ``scores`` / ``counts`` + ``publish`` in one call
        (``gmm_evaluate_publish``): the Contacts kernel stores the rows into every rank's table as
        the individuals finish."""
        from . import _lib
        C = self._C
        ps, pc = C.c_void_p(), C.c_void_p()
        p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        _lib.check(self._L.gmm_evaluate_publish(
            self.engine._h, int(scores.shape[0]), p(torsions), p(xforms), p(rot_alleles), None, p(scores), p(counts),
            int(row0), C.byref(ps), C.byref(pc), self.engine._stream()))
        return self._views(ps, pc)

    def state(self):
        """Synchronise; ``(epochs completed, timed_out, scores view, counts view)`` of the last epoch."""
        from . import _lib
        C = self._C
        epoch, err, ps, pc = C.c_uint(0), C.c_int(0), C.c_void_p(), C.c_void_p()
        _lib.check(self._L.gmm_exchange_state(self.engine._h, C.byref(epoch), C.byref(err), C.byref(ps), C.byref(pc)))
        s, c = self._views(ps, pc)
        return int(epoch.value), bool(err.value), s, c

    def close(self):
        self._L.gmm_exchange_destroy(self.engine._h)


class ShardedEvaluator(object):
    """Population sharded by individual over the ranks of one node, device-resident alleles:
    rank r evaluates rows ``[bounds[r], bounds[r + 1])`` of the population and every rank ends up
    with the whole ``[P, 2]`` fitness table (rank 0 runs NSGA-II selection,
    gaudi/algorithms.py:160).  ``exchange``: ``"peer"`` (default: stores through peer memory,
    :class:`FitnessExchange`) or ``"nccl"`` (``all_gather_into_tensor``; the cross-check)."""

    def __init__(self, engine, population, group=None, exchange="peer"):
        self.engine, self.group, self.population = engine, group, int(population)
        self.world, self.rank = _world(group)
        self.bounds = shard_bounds(self.population, self.world)
        self.lo, self.hi = self.bounds[self.rank], self.bounds[self.rank + 1]
        self.max_shard = max_shard(self.population, self.world)
        dev = engine.torch_device
        self.scores = torch.zeros((self.max_shard, 2), dtype=torch.float64, device=dev)
        self.counts = torch.zeros((self.max_shard, 2), dtype=torch.int32, device=dev)
        engine.reserve(self.max_shard)
        self.kind = exchange if self.world > 1 else "none"
        self.exchange = FitnessExchange(engine, self.population, group) if self.kind == "peer" else None
        if self.kind == "nccl":
            self._gs = torch.empty((self.world * self.max_shard, 2), dtype=torch.float64, device=dev)
            self._gc = torch.empty((self.world * self.max_shard, 2), dtype=torch.int32, device=dev)

    def shard(self, a):
        """This rank's rows of a population-sized array (or None)."""
        return None if a is None else a[self.lo:self.hi]

    def evaluate(self, torsions=None, xforms=None, rot_alleles=None):
        """``torsions / xforms / rot_alleles``: THIS rank's rows, on the device.  Returns
        ``(scores [P, 2] fp64, counts [P, 2] int32)`` with every rank's rows, on every rank."""
        n = self.hi - self.lo
        s, c = self.scores[:n], self.counts[:n]
        if self.kind == "peer":
            eng = self.engine
            t, x, r = (eng._dev(torsions, torch.float64), eng._dev(xforms, torch.float64),
                       eng._dev(rot_alleles, torch.float64))
            gs, gc = self.exchange.evaluate_publish(t, x, r, s, c, self.lo)
            return gs[:self.population], gc[:self.population]
        self.engine.evaluate(torsions, xforms, rot_alleles, scores_out=s, counts_out=c)
        if self.kind == "none":
            return s, c
        _all_gather_into(self._gs, self.scores, self.group)
        _all_gather_into(self._gc, self.counts, self.group)
        m, b = self.max_shard, self.bounds
        if all(b[r + 1] - b[r] == m for r in range(self.world)):
            return self._gs, self._gc
        return (torch.cat([self._gs[r * m:r * m + b[r + 1] - b[r]] for r in range(self.world)]),
                torch.cat([self._gc[r * m:r * m + b[r + 1] - b[r]] for r in range(self.world)]))

    def close(self):
        if self.exchange is not None:
            self.exchange.close()


# ---------------------------------------------------------------------------------------------
# Host-resident populations on several GPUs: one page-locked table shared by the ranks
# ---------------------------------------------------------------------------------------------
class HostSharded(object):
    """The GA lives in ONE process (rank 0: DEAP-style selection, variation and niching are
    sequential host code, gaudi/algorithms.py:54-202); its allele and fitness tables live in a
    shared-memory segment that every rank maps and page-locks (``cudaHostRegister``).  A step is
    then: rank 0 raises the epoch; every rank copies ITS rows host-to-device over its own PCIe
    link, evaluates them, and copies its results device-to-host into the same table; rank 0 waits
    for the ranks' done marks and reads the whole table.  No allele or score crosses NVLink, and
    the upload runs on ``world`` PCIe links in parallel instead of through rank 0's.

    ``spec``: ordered ``name -> (trailing shape, numpy dtype)``; ``self.arrays[name]`` is the
    ``[max_rows, ...]`` table.  ``run(n_rows, fn)`` is collective: every rank calls it (ranks other
    than 0 usually through ``serve``), and ``fn(lo, hi)`` runs on each rank with a non-empty block.
    The process group is used at construction only (segment name, barrier)."""

    HEADER = 64     # int64 slots: 0 epoch, 1 n_rows, 2 shutdown, 8 + r: last epoch rank r has finished

    def __init__(self, spec, max_rows, group=None, register=True, timeout=120.0):
        import os
        import uuid
        import numpy as np
        self.group, self.max_rows, self.timeout = group, int(max_rows), float(timeout)
        self.world, self.rank = _world(group)
        offsets, at = {}, self.HEADER * 8
        for name, (shape, dtype) in spec.items():
            nbytes = int(np.prod((self.max_rows,) + tuple(shape))) * np.dtype(dtype).itemsize
            offsets[name] = at
            at += (nbytes + 255) // 256 * 256
        self.nbytes = max(at, 4096)
        path = [None]
        if self.rank == 0:
            import tempfile
            base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else tempfile.gettempdir()
            path[0] = os.path.join(base, "gaudimm_b200_%s" % uuid.uuid4().hex)
            with open(path[0], "wb") as fh:
                fh.truncate(self.nbytes)
        if self.world > 1:
            dist.broadcast_object_list(path, src=0, group=group)
        self.path = path[0]
        self._map = np.memmap(self.path, dtype=np.uint8, mode="r+", shape=(self.nbytes,))
        self.header = self._map[:self.HEADER * 8].view(np.int64)
        self.arrays = {}
        for name, (shape, dtype) in spec.items():
            n = int(np.prod((self.max_rows,) + tuple(shape))) * np.dtype(dtype).itemsize
            self.arrays[name] = self._map[offsets[name]:offsets[name] + n].view(dtype).reshape((self.max_rows,) + tuple(shape))
        self._registered = False
        if register and torch.cuda.is_available():
            rc = torch.cuda.cudart().cudaHostRegister(self._map.ctypes.data, self.nbytes, 0)
            if int(rc) != 0:
                raise RuntimeError("cudaHostRegister of the shared population table failed with code %d" % int(rc))
            self._registered = True
        self._seen = 0
        if self.world > 1:
            dist.barrier(group=group)
        if self.rank == 0:
            os.unlink(self.path)          # the mappings keep the segment alive; nothing is left behind on a crash

    def _wait(self, cond, what):
        """Spin on a header slot: a few milliseconds flat out (a step is a fraction of a millisecond),
        then with short sleeps, so that ranks idling between generations do not burn a core each."""
        import time
        t0 = time.perf_counter()
        spins = 0
        while not cond():
            spins += 1
            if (spins & 0xff) == 0:
                waited = time.perf_counter() - t0
                if waited > self.timeout:
                    raise RuntimeError("HostSharded: rank %d waited more than %.0f s for %s" % (self.rank, self.timeout, what))
                if waited > 5e-3:
                    time.sleep(2e-4)

    def run(self, n_rows, fn):
        """Collective.  Rank 0's ``n_rows`` counts; returns the row bounds used."""
        h = self.header
        if self.rank == 0:
            if n_rows > self.max_rows:
                raise ValueError("population of %d rows exceeds the table's %d" % (n_rows, self.max_rows))
            h[1] = int(n_rows)
            h[0] = self._seen + 1
        self._wait(lambda: h[0] > self._seen or h[2] != 0, "rank 0 to start a step")
        if h[2] != 0:
            return None
        self._seen = int(h[0])
        b = shard_bounds(int(h[1]), self.world)
        lo, hi = b[self.rank], b[self.rank + 1]
        if hi > lo:
            fn(lo, hi)
        h[8 + self.rank] = self._seen
        if self.rank == 0:
            for r in range(1, self.world):
                self._wait(lambda r=r: h[8 + r] >= self._seen, "rank %d to finish its block" % r)
        return b

    def serve(self, fn):
        """Ranks other than 0: take part in rank 0's steps until it calls ``shutdown``."""
        while self.run(0, fn) is not None:
            pass

    def shutdown(self):
        if self.rank == 0:
            self.header[2] = 1

    def close(self):
        if self._registered:
            torch.cuda.synchronize()
            torch.cuda.cudart().cudaHostUnregister(self._map.ctypes.data)
            self._registered = False


class HostShardedEvaluator(object):
    """``gmm_evaluate_host`` on several GPUs for a population that lives in rank 0's host memory
    (the ``toolbox.map(toolbox.evaluate, population)`` seam, gaudi/algorithms.py:99-102): the allele
    and fitness tables are the page-locked shared tables of :class:`HostSharded`; each rank's
    ``gmm_evaluate_host`` copies its rows straight out of and into them.

    Rank 0 fills ``tors / xforms / rot`` (numpy views, ``[max_pop, ...]``) and calls
    ``evaluate(P)``; the other ranks call ``serve()`` (or ``evaluate()`` once per step).  After the
    call rank 0 reads ``scores[:P]`` / ``counts[:P]`` (and ``xyz[:P]`` when ``want_xyz``)."""

    def __init__(self, engine, max_pop, group=None, want_xyz=False):
        from collections import OrderedDict
        import numpy as np
        m = engine.mobile
        self.engine, self.want_xyz = engine, bool(want_xyz)
        spec = OrderedDict()
        spec["tors"] = ((m.n_allele_torsions,), np.float64)
        spec["xforms"] = ((12,), np.float64)
        spec["rot"] = ((m.n_rotamer_residues,), np.float64)
        spec["scores"] = ((2,), np.float64)
        spec["counts"] = ((2,), np.int32)
        if want_xyz:
            spec["xyz"] = ((m.n_atoms, 3), np.float32)
        self.table = HostSharded(spec, max_pop, group)
        for name, arr in self.table.arrays.items():
            setattr(self, name, arr)
        self.world, self.rank = self.table.world, self.table.rank
        engine.reserve(max_shard(int(max_pop), self.world))
        self.use_xforms = True

    def _block(self, lo, hi):
        a = self.table.arrays
        self.engine.evaluate_host(a["tors"][lo:hi], a["xforms"][lo:hi] if self.use_xforms else None,
                                  a["rot"][lo:hi] if a["rot"].shape[1] else None,
                                  scores_out=a["scores"][lo:hi], counts_out=a["counts"][lo:hi],
                                  xyz_out=a["xyz"][lo:hi] if self.want_xyz else None)

    def evaluate(self, n_rows=0):
        return self.table.run(n_rows, self._block)

    def serve(self):
        self.table.serve(self._block)

    def shutdown(self):
        self.table.shutdown()

    def close(self):
        self.table.close()
