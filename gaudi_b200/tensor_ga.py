"""Population-as-tensor GA loop ("next" row 1).

``gaudi_b200.algorithms`` keeps one Python object per individual, like the reference; that is
fine for the populations GaudiMM documents (100 - 1,000, docs/faq.rst:48-50) but the per-object
work (deepcopy per clone, gaudi/base.py:289-299; O((mu+lambda)^2) Python ``similar`` calls,
gaudi/algorithms.py:139-155; DEAP's selNSGA2) becomes the bottleneck long before the GPU does.
Here the whole population is a set of arrays:

    torsions  (N, T)      degrees, rounded to the Torsion gene's precision
    trans     (N, 3)      translation column of the Search allele's first matrix
    rot       (N, 3, 3)   rotation block of the Search allele's second matrix
    rotamers  (N, n_res)  Rotamers alleles in [0, 1)
    fitness   (N, n_obj)  raw objective values (weights applied on comparison, as DEAP does)

and every step of gaudi/algorithms.py:54-202 is an array operation: variation (the same
operators with the same parameters as the gene plugins: SBX + polynomial mutation / random reset
for Torsion, torsion.py:173-195; xform interpolation for Search, search.py:222-246; two-point
crossover and per-allele resampling for Rotamers, rotamers.py:159-163), one ``gmm_evaluate`` per
generation, one ``gmm_rmsd_matrix`` per generation feeding an exact re-statement of the greedy
niching pass, and NSGA-II selection.  Host side is numpy; the scoring is the CUDA path.
"""
from __future__ import annotations

import time

import numpy as np


# ---- NSGA-II on arrays ---------------------------------------------------------------------
_fronts_backend = None      # set by launch(): Engine.nondominated_fronts (GPU) for three or more objectives


def set_fronts_backend(fn):
    """``fn(w) -> fronts`` used for more than two objectives (``gmm_nondominated_fronts`` through an
    Engine); ``None`` restores the host restatement (O(N^2) numpy, small populations only)."""
    global _fronts_backend
    _fronts_backend = fn


def nondominated_fronts(w):
    """Front index of every row of ``w`` (N, M) under maximisation of each column (``wvalues``).
    Equal rows share a front, as in DEAP's sortNondominated (which groups by fitness).  Two
    objectives: O(N log N) on the host.  More: the GPU sort when a backend is set."""
    n = w.shape[0]
    if w.shape[1] != 2 and _fronts_backend is not None and n > 256:
        return np.asarray(_fronts_backend(w), dtype=np.int64)
    if w.shape[1] == 2 and n > 0:
        # O(N log N) for two objectives (gmm_fronts2_host): rows visited by first objective
        # descending, ties by second descending, so every dominator of a row comes before it; front
        # f keeps the largest second objective it holds, and a row joins the first front whose best
        # is strictly below its own second objective.
        import ctypes as C
        from . import _lib
        order = np.lexsort((-w[:, 1], -w[:, 0]))
        ws = np.ascontiguousarray(w[order], dtype=np.float64)
        lev = np.empty(n, dtype=np.int32)
        _lib.check(_lib.lib().gmm_fronts2_host(int(n), ws.ctypes.data_as(C.c_void_p), lev.ctypes.data_as(C.c_void_p)))
        front = np.empty(n, dtype=np.int64)
        front[order] = lev
        return front
    front = np.full(n, -1, dtype=np.int64)
    # any other number of objectives, on the host: peel the non-dominated layer of the unique rows
    # (O(N^2) per layer: small populations only; launch() installs the device sort)
    uniq, inv = np.unique(w, axis=0, return_inverse=True)
    inv = inv.reshape(-1)
    u = uniq.shape[0]
    ufront = np.full(u, -1, dtype=np.int64)
    remaining = np.arange(u)
    level = 0
    while len(remaining):
        sub = uniq[remaining]
        nd = np.ones(len(sub), dtype=bool)
        for s in range(0, len(sub), 512):          # chunked O(n^2) dominance test
            blk = sub[s:s + 512]
            ge = (sub[None, :, :] >= blk[:, None, :]).all(axis=2)
            gt = (sub[None, :, :] > blk[:, None, :]).any(axis=2)
            nd[s:s + 512] = ~(ge & gt).any(axis=1)
        ufront[remaining[nd]] = level
        remaining = remaining[~nd]
        level += 1
    front[:] = ufront[inv]
    return front


def crowding_distance(values):
    """DEAP's assignCrowdingDist on raw ``values`` (n, M) of one front."""
    n, m = values.shape
    dist = np.zeros(n)
    if n == 0:
        return dist
    for i in range(m):
        order = np.argsort(values[:, i], kind="stable")
        v = values[order, i]
        dist[order[0]] = np.inf
        dist[order[-1]] = np.inf
        if v[-1] == v[0]:
            continue
        norm = m * float(v[-1] - v[0])
        dist[order[1:-1]] += (v[2:] - v[:-2]) / norm
    return dist


def sel_nsga2(values, weights, k):
    """Indices selected by NSGA-II (``deap.tools.selNSGA2``): whole fronts first, the last one
    by decreasing crowding distance."""
    w = values * np.asarray(weights)[None, :]
    front = nondominated_fronts(w)
    n = len(front)
    if k >= n:
        return np.argsort(front, kind="stable")
    order = np.argsort(front, kind="stable")           # whole fronts first, original order inside a front
    counts = np.bincount(front)
    ends = np.cumsum(counts)
    last = int(np.searchsorted(ends, k, side="left"))  # the front that crosses k (or ends exactly at it)
    start = int(ends[last - 1]) if last > 0 else 0
    chosen = order[:start]
    if start < k:
        idx = order[start:ends[last]]
        if ends[last] == k:
            chosen = order[:k]
        else:
            cd = crowding_distance(values[idx])
            pick = np.argsort(-cd, kind="stable")[:k - start]
            chosen = np.concatenate([chosen, idx[pick]])
    return np.asarray(chosen, dtype=np.int64)


# ---- greedy niching on a similarity bitmask ------------------------------------------------
def lex_rank(w):
    """Dense rank of the rows of ``w`` in lexicographic order (what tuple comparison of DEAP's
    ``wvalues`` does, used by ``i.fitness <= n_i.fitness`` at gaudi/algorithms.py:148)."""
    order = np.lexsort(tuple(w[:, c] for c in range(w.shape[1] - 1, -1, -1)))
    ranks = np.empty(len(w), dtype=np.int64)
    sw = w[order]
    new = np.ones(len(w), dtype=bool)
    new[1:] = (sw[1:] != sw[:-1]).any(axis=1)
    ranks[order] = np.cumsum(new) - 1
    return ranks


def niching(sim, w):
    """gaudi/algorithms.py:139-155 on a boolean similarity matrix ``sim`` (N, N) and wvalues
    ``w`` (N, M), candidates visited in index order.  The reference's inner loop breaks (drops the
    candidate, deletes nobody) as soon as a similar kept individual is at least as fit, and only
    when there is none evicts every similar kept one; hence per candidate: drop it if any similar
    kept individual has rank >= its own, else evict them all and keep it.  Returns kept indices in
    the reference's order."""
    if isinstance(sim, tuple):                 # (ptr, idx) partner lists from the GPU compaction
        from .engine import niching_lists
        return niching_lists(sim[0], sim[1], lex_rank(w))
    if sim.dtype != np.bool_:
        return niching_bits(sim, w)
    n = sim.shape[0]
    rank = lex_rank(w)
    kept = np.zeros(n, dtype=bool)
    for i in range(n):
        hits = np.flatnonzero(sim[i] & kept)
        if len(hits):
            if (rank[hits] >= rank[i]).any():
                continue
            kept[hits] = False
        kept[i] = True
    return np.flatnonzero(kept)


def niching_bits(words, w):
    """``niching`` on the packed bitmask ``gmm_rmsd_matrix`` writes: ``words`` (N, ceil(N/32))
    uint32/int32, bit ``j & 31`` of word ``j >> 5`` of row ``i`` set when i and j are similar.  A
    population of 131,072 candidates is a 2 GiB bitmask; the boolean matrix would be 17 GB."""
    words = np.ascontiguousarray(words).view(np.uint32)
    n = words.shape[0]
    rank = lex_rank(w)
    kept = np.zeros(n, dtype=bool)
    kept_words = np.zeros(words.shape[1], dtype=np.uint32)
    for i in range(n):
        both = words[i] & kept_words
        nz = np.flatnonzero(both)
        if len(nz):
            hits = []
            for k in nz.tolist():
                v = int(both[k])
                while v:
                    hits.append(32 * k + ((v & -v).bit_length() - 1))
                    v &= v - 1
            hits = np.asarray(hits, dtype=np.int64)
            if (rank[hits] >= rank[i]).any():
                continue
            kept[hits] = False
            for h in hits.tolist():
                kept_words[h >> 5] &= np.uint32(~(1 << (h & 31)) & 0xFFFFFFFF)
        kept[i] = True
        kept_words[i >> 5] |= np.uint32(1 << (i & 31))
    return np.flatnonzero(kept)


# ---- rotations -------------------------------------------------------------------------------
def random_rotations(n, rng):
    q = rng.normal(size=(n, 4))
    return quat_to_matrix(q / np.linalg.norm(q, axis=1, keepdims=True))


def quat_to_matrix(q):
    w, x, y, z = q.T
    r = np.empty((len(q), 3, 3))
    r[:, 0, 0] = 1 - 2 * (y * y + z * z); r[:, 0, 1] = 2 * (x * y - z * w); r[:, 0, 2] = 2 * (x * z + y * w)
    r[:, 1, 0] = 2 * (x * y + z * w); r[:, 1, 1] = 1 - 2 * (x * x + z * z); r[:, 1, 2] = 2 * (y * z - x * w)
    r[:, 2, 0] = 2 * (x * z - y * w); r[:, 2, 1] = 2 * (y * z + x * w); r[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return r


def matrix_to_quat(r):
    n = len(r)
    q = np.empty((n, 4))
    tr = r[:, 0, 0] + r[:, 1, 1] + r[:, 2, 2]
    q[:, 0] = np.sqrt(np.maximum(0.0, 1 + tr)) / 2
    q[:, 1] = np.copysign(np.sqrt(np.maximum(0.0, 1 + r[:, 0, 0] - r[:, 1, 1] - r[:, 2, 2])) / 2, r[:, 2, 1] - r[:, 1, 2])
    q[:, 2] = np.copysign(np.sqrt(np.maximum(0.0, 1 - r[:, 0, 0] + r[:, 1, 1] - r[:, 2, 2])) / 2, r[:, 0, 2] - r[:, 2, 0])
    q[:, 3] = np.copysign(np.sqrt(np.maximum(0.0, 1 - r[:, 0, 0] - r[:, 1, 1] + r[:, 2, 2])) / 2, r[:, 1, 0] - r[:, 0, 1])
    return q / np.linalg.norm(q, axis=1, keepdims=True)


def slerp(r1, r2, frac):
    """Rotation a fraction of the way from r1 to r2 (batched), the rotational part of the
    xform interpolation used at search.py:222-246."""
    q1, q2 = matrix_to_quat(r1), matrix_to_quat(r2)
    d = (q1 * q2).sum(axis=1)
    q2 = np.where(d[:, None] < 0, -q2, q2)
    d = np.abs(d)
    th = np.arccos(np.clip(d, -1, 1))
    s = np.sin(th)
    lin = s < 1e-6
    a = np.where(lin, 1 - frac, np.sin((1 - frac) * th) / np.where(lin, 1, s))
    b = np.where(lin, frac, np.sin(frac * th) / np.where(lin, 1, s))
    q = a[:, None] * q1 + b[:, None] * q2
    return quat_to_matrix(q / np.linalg.norm(q, axis=1, keepdims=True))


# ---- the loop ----------------------------------------------------------------------------------
class TensorGA(object):
    """(mu + lambda) NSGA-II GA over array alleles, scored by a ``gaudi_b200.base.BatchEvaluator``.

    ``spec`` carries what the gene plugins would: ``n_torsions, flexibility, precision, cx_eta,
    mut_eta, mut_indpb, search_center, search_radius, rotate, interpolation, n_rot_res``.
    ``evaluate(torsions, xforms, rotamers) -> (scores (P, n_obj) rounded, coords (P, A, 3) float32)``.
    ``similarity(coords, threshold) -> (N, N) bool``."""

    def __init__(self, spec, weights, evaluate, similarity, threshold=2.5, seed=None):
        self.spec = dict(spec)
        self.weights = np.asarray(weights, dtype=np.float64)
        self.evaluate = evaluate
        self.similarity = similarity
        self.threshold = threshold
        self.rng = np.random.default_rng(seed)

    # population construction ------------------------------------------------------------------
    def random_population(self, n):
        s, rng = self.spec, self.rng
        half = 0.5 * s["flexibility"]
        pop = {
            "torsions": np.round(rng.uniform(-half, half, size=(n, s["n_torsions"])), s["precision"]),
            "trans": self._random_translations(n),
            "rot": random_rotations(n, rng) if s.get("rotate", True) else np.tile(np.eye(3), (n, 1, 1)),
            "rotamers": rng.random((n, s.get("n_rot_res", 0))),
            "fitness": np.full((n, len(self.weights)), np.nan),
            "coords": None,
        }
        return pop

    def _random_translations(self, n):
        c, r = np.asarray(self.spec["search_center"], float), float(self.spec["search_radius"])
        out = np.empty((0, 3))
        while len(out) < n:                                    # search.py:311-335
            p = self.rng.uniform(-r, r, size=(2 * n + 16, 3))
            out = np.concatenate([out, p[(p * p).sum(axis=1) <= r * r]])
        return out[:n] + c

    @staticmethod
    def xforms(pop, precision=0):
        """(N, 12) products ``T . R`` for ``gmm_express``; with ``Search.precision > 0`` both allele
        matrices are rounded to that many decimals first (search.py:207-210)."""
        rot, trans = pop["rot"], pop["trans"]
        prec = int(precision)
        if prec > 0:
            rot, trans = np.around(rot, prec), np.around(trans, prec)
        return np.concatenate([rot, trans[:, :, None]], axis=2).reshape(-1, 12)

    @staticmethod
    def take(pop, idx):
        return {k: (v[idx] if v is not None else None) for k, v in pop.items()}

    @staticmethod
    def concat(a, b):
        return {k: (np.concatenate([a[k], b[k]]) if a[k] is not None else None) for k in a}

    def score(self, pop, mask=None):
        idx = np.flatnonzero(np.isnan(pop["fitness"][:, 0]) if mask is None else mask)
        if len(idx) == 0:
            return 0
        sub = self.take(pop, idx)
        scores, coords = self.evaluate(sub["torsions"], self.xforms(sub, self.spec.get("search_precision", 0)), sub["rotamers"])
        pop["fitness"][idx] = scores
        if pop["coords"] is None:
            pop["coords"] = np.zeros((len(pop["fitness"]),) + coords.shape[1:], dtype=np.float32)
        pop["coords"][idx] = coords
        return len(idx)

    # variation (deap.algorithms.varOr over arrays) ------------------------------------------------
    def var_or(self, pop, lambda_, cxpb, mutpb):
        s, rng = self.spec, self.rng
        n = len(pop["fitness"])
        op = rng.random(lambda_)
        p1 = rng.integers(0, n, size=lambda_)
        p2 = (p1 + rng.integers(1, max(n, 2), size=lambda_)) % n          # a different parent
        child = self.take(pop, p1)
        child = {k: (v.copy() if v is not None else None) for k, v in child.items()}
        cx = op < cxpb
        mu = (~cx) & (op < cxpb + mutpb)
        half = 0.5 * s["flexibility"]
        if cx.any():
            k = np.flatnonzero(cx)
            a, b = pop["torsions"][p1[k]], pop["torsions"][p2[k]]
            child["torsions"][k] = np.round(self._sbx(a, b, s["cx_eta"], -half, half), s["precision"])
            # Search.mate: first child keeps its translation, takes the half-way rotation
            child["rot"][k] = slerp(pop["rot"][p1[k]], pop["rot"][p2[k]], 0.5)
            if child["rotamers"].shape[1] >= 2:                       # cxTwoPoint, first child
                m = child["rotamers"].shape[1]
                lo = rng.integers(1, m + 1, size=len(k)); hi = rng.integers(1, m, size=len(k))
                hi = np.where(hi >= lo, hi + 1, hi)
                lo, hi = np.minimum(lo, hi), np.maximum(lo, hi)
                cols = np.arange(m)[None, :]
                swap = (cols >= lo[:, None]) & (cols < hi[:, None])
                child["rotamers"][k] = np.where(swap, pop["rotamers"][p2[k]], child["rotamers"][k])
            child["fitness"][k] = np.nan
        if mu.any():
            k = np.flatnonzero(mu)
            t = child["torsions"][k]
            poly = rng.random(len(k)) < 0.5                              # torsion.py:185
            t_poly = np.round(self._poly(t, s["mut_eta"], -half, half, s["mut_indpb"]), s["precision"])
            t_new = np.round(rng.uniform(-half, half, size=t.shape), s["precision"])
            child["torsions"][k] = np.where(poly[:, None], t_poly, t_new)
            hit = rng.random(len(k)) < s.get("search_indpb", s["mut_indpb"])   # search.py:236
            if hit.any():
                kk = k[hit]
                f = s.get("interpolation", 0.5)
                tgt_r = random_rotations(len(kk), rng) if s.get("rotate", True) else child["rot"][kk]
                tgt_t = self._random_translations(len(kk))
                child["rot"][kk] = slerp(child["rot"][kk], tgt_r, f)
                child["trans"][kk] = child["trans"][kk] + f * (tgt_t - child["trans"][kk])
            if child["rotamers"].shape[1]:
                r = child["rotamers"][k]
                resample = rng.random(r.shape) < s.get("rotamers_indpb", s["mut_indpb"])   # rotamers.py:163
                child["rotamers"][k] = np.where(resample, rng.random(r.shape), r)
            child["fitness"][k] = np.nan
        return child

    def _sbx(self, a, b, eta, low, up):
        """First child of cxSimulatedBinaryBounded, vectorised over (individual, locus)."""
        rng = self.rng
        x1, x2 = np.minimum(a, b), np.maximum(a, b)
        active = (rng.random(a.shape) <= 0.5) & (np.abs(a - b) > 1e-14)
        diff = np.where(active, x2 - x1, 1.0)
        rand = rng.random(a.shape)

        def betaq(beta):
            alpha = 2.0 - beta ** -(eta + 1)
            return np.where(rand <= 1.0 / alpha, (rand * alpha) ** (1.0 / (eta + 1)),
                            (1.0 / (2.0 - rand * alpha)) ** (1.0 / (eta + 1)))
        c1 = 0.5 * (x1 + x2 - betaq(1.0 + 2.0 * (x1 - low) / diff) * (x2 - x1))
        c2 = 0.5 * (x1 + x2 + betaq(1.0 + 2.0 * (up - x2) / diff) * (x2 - x1))
        c1, c2 = np.clip(c1, low, up), np.clip(c2, low, up)
        first = np.where(rng.random(a.shape) <= 0.5, c2, c1)
        return np.where(active, first, a)

    def _poly(self, x, eta, low, up, indpb):
        rng = self.rng
        hit = rng.random(x.shape) <= indpb
        d1, d2 = (x - low) / (up - low), (up - x) / (up - low)
        rand = rng.random(x.shape)
        mp = 1.0 / (eta + 1.0)
        lo_branch = (2.0 * rand + (1.0 - 2.0 * rand) * (1.0 - d1) ** (eta + 1)) ** mp - 1.0
        hi_branch = 1.0 - (2.0 * (1.0 - rand) + 2.0 * (rand - 0.5) * (1.0 - d2) ** (eta + 1)) ** mp
        dq = np.where(rand < 0.5, lo_branch, hi_branch)
        return np.where(hit, np.clip(x + dq * (up - low), low, up), x)

    # the generational loop (gaudi/algorithms.py:54-202) -------------------------------------------
    def run(self, population, mu, lambda_, cxpb, mutpb, ngen):
        log = []
        t0 = time.time()
        nev = self.score(population)
        order = sel_nsga2(population["fitness"], self.weights, len(population["fitness"]))
        population = self.take(population, order)
        log.append({"gen": 0, "nevals": nev, "speed": nev / max(time.time() - t0, 1e-9)})
        for gen in range(1, ngen + 1):
            t1 = time.time()
            offspring = self.var_or(population, lambda_, cxpb, mutpb)
            t2 = time.time()
            # children made by plain reproduction carry their parent's fitness and coordinates
            nev = self.score(offspring)
            t3 = time.time()
            cand = self.concat(population, offspring)
            w = cand["fitness"] * self.weights[None, :]
            sim = self.similarity(cand["coords"], self.threshold)
            t4 = time.time()
            kept = niching(sim, w)
            if len(kept) < mu:                                   # algorithms.py:156-158
                extra = self.rng.integers(0, len(w), size=mu - len(kept))
                kept = np.concatenate([kept, extra])
            new_pop = self.take(cand, kept)
            t5 = time.time()
            sel = sel_nsga2(new_pop["fitness"], self.weights, mu)
            population = self.take(new_pop, sel)
            t6 = time.time()
            log.append({"gen": gen, "nevals": nev, "speed": nev / max(t3 - t2, 1e-9),
                        "seconds": {"variation": t2 - t1, "evaluate": t3 - t2, "similarity": t4 - t3,
                                    "niching": t5 - t4, "select": t6 - t5, "generation": t6 - t1}})
        return population, log


def py_round(a, ndigits):
    """Element-wise Python ``round`` (what gaudi/base.py:361 applies), not numpy's scaled rint."""
    flat = np.asarray(a, dtype=np.float64).ravel()
    return np.array([round(float(v), ndigits) for v in flat]).reshape(np.shape(a))


def launch(cfg, device=0, seed=None, group=None):
    """``gaudi_b200.algorithms.launch`` with the population held as arrays.  Same job description;
    one template individual is instantiated to discover the topology, then the GA never touches
    per-individual Python objects.  Returns ``(population dict, logbook, BatchEvaluator)``.  Under a
    one-process-per-GPU launch rank 0 runs the GA, the other ranks evaluate their blocks and return
    ``(None, None, batch)``, as in ``gaudi_b200.algorithms.launch``."""
    import random
    from . import base, parallel, plugin
    if seed is not None:
        random.seed(seed)
    plugin.import_plugins(*cfg.genes)
    plugin.import_plugins(*cfg.objectives)
    template = base.MolecularIndividual(cfg)
    env = base.Environment(cfg, device=device, group=group)
    cfg._environment = env
    batch = env.attach(template)
    n_pop = cfg.ga.population
    batch.enable_sharding(n_pop + int(cfg.ga.lambda_ * n_pop))
    if (parallel._world(group)[1] if group is not False else 0) != 0:
        batch.serve()
        return None, None, batch
    tors = template.genes.get(batch.torsion_name) if batch.torsion_name else None
    search = template.genes.get(batch.search_name) if batch.search_name else None
    objectives = list(env.objectives.values())
    spec = {
        "n_torsions": batch.n_torsions,
        "flexibility": tors.flexibility if tors is not None else 360.0,
        "precision": tors.precision if tors is not None else 1,
        # the gene instances carry the job-wide ga.* values unless their own YAML block overrides them
        # (plugin.load_plugins passes both; gaudi/genes/__init__.py:70-80)
        "cx_eta": tors.cxeta if tors is not None else float(cfg.ga.cx_eta),
        "mut_eta": tors.mteta if tors is not None else float(cfg.ga.mut_eta),
        "mut_indpb": tors.indpb if tors is not None else float(cfg.ga.mut_indpb),
        "search_indpb": search.indpb if search is not None else float(cfg.ga.mut_indpb),
        "rotamers_indpb": (template.genes[batch.rotamer_name].indpb if batch.rotamer_name
                           else float(cfg.ga.mut_indpb)),
        "search_precision": search.precision if search is not None else 0,
        "search_center": search.center if search is not None else batch.centre,
        "search_radius": search.radius if search is not None else 0.0,
        "rotate": search.rotate if search is not None else False,
        "interpolation": search.interpolation if search is not None else 0.5,
        "n_rot_res": batch.n_rot_res,
    }

    def evaluate(t, xf, r):
        # every GPU of the job when launched one process per GPU (BatchEvaluator.score_population)
        raw, xyz = batch.score_population(t, xf if search is not None else None, r if batch.n_rot_res else None)
        out = np.empty((len(t), len(objectives)))
        for i, o in enumerate(objectives):
            out[:, i] = py_round(raw[:, i], o.precision)
        return out, np.ascontiguousarray(xyz[:, batch.lig_slice])

    def similarity(coords, threshold):
        # the N x N bitmask stays on the device; only the (sparse) partner lists come back
        eng = batch.engines[0][0]
        _, bits = eng.rmsd_matrix(coords, threshold=threshold, values=False)
        return eng.similar_lists(bits, len(coords))

    set_fronts_backend(batch.engines[0][0].nondominated_fronts)
    ga = TensorGA(spec, cfg.weights, evaluate, similarity, threshold=cfg.similarity.args[1], seed=seed)
    n = cfg.ga.population
    pop = ga.random_population(n)
    mu, lambda_ = int(cfg.ga.mu * n), int(cfg.ga.lambda_ * n)
    pop, log = ga.run(pop, mu, lambda_, cfg.ga.cx_pb, cfg.ga.mut_pb, cfg.ga.generations)
    batch.shutdown()
    return pop, log, batch
