"""
Static model of a GaudiMM job as the CUDA path consumes it.

The reference keeps topology inside live ``chimera.Molecule`` objects and rediscovers it
on every evaluation (``BondRot`` objects at gaudi/genes/torsion.py:245-307, 4-level bond
BFS inside ``DetectClash`` at gaudi/objectives/contacts.py:142-145).  Here everything that
does not depend on the alleles is flattened once, on the host, into plain arrays:

* ``MobileModel``   – atoms whose coordinates depend on the genome (the Search/Torsion
  target ligand and, with a Rotamers gene, the flexible side chains);
* ``ReceptorModel`` – every other atom of every Molecule gene (fixed coordinates);
* ``ContactParams`` – the Contacts objective's settings (contacts.py:101-121).

These arrays are exactly what ``include/gaudimm_b200.h`` takes.  This is input
preparation (topology walks, run once per job); no scoring arithmetic happens here.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

# per-atom flag bits (uint8) ----------------------------------------------------------
F_HYDROPHOBIC = 1   # element in Contacts.hydrophobic_elements (contacts.py:103)
F_DONOR = 2         # H-bond donor   (DetectClash hbondAllowance, contacts.py:144)
F_ACCEPTOR = 4      # H-bond acceptor
F_SEED = 8          # zone seed: atom of a `probes` molecule or CB of a Rotamers residue
F_XFORM = 16        # mobile atom that the Search gene's xform applies to


def adjacency(n_atoms, bonds):
    adj = [[] for _ in range(n_atoms)]
    for a, b in np.asarray(bonds, dtype=np.int64).reshape(-1, 2):
        adj[int(a)].append(int(b))
        adj[int(b)].append(int(a))
    return adj


def within_bonds(adj, src, depth):
    """Atoms reachable from ``src`` in at most ``depth`` bonds (``src`` included)."""
    seen = {src}
    frontier = [src]
    for _ in range(depth):
        nxt = []
        for a in frontier:
            for b in adj[a]:
                if b not in seen:
                    seen.add(b)
                    nxt.append(b)
        frontier = nxt
        if not frontier:
            break
    return seen


def side_of_bond(adj, fixed, moving):
    """Atoms on the ``moving`` side of the bond fixed-moving (``moving`` included).
    Raises ValueError if the bond is in a cycle (``chimera.BondRot`` refuses those,
    torsion.py:299-304)."""
    seen = {moving}
    stack = [moving]
    while stack:
        a = stack.pop()
        for b in adj[a]:
            if a == moving and b == fixed:
                continue
            if b == fixed:
                raise ValueError("bond %d-%d is in a cycle" % (fixed, moving))
            if b not in seen:
                seen.add(b)
                stack.append(b)
    return seen


def graph_distances(adj, src):
    dist = {src: 0}
    frontier = [src]
    while frontier:
        nxt = []
        for a in frontier:
            for b in adj[a]:
                if b not in dist:
                    dist[b] = dist[a] + 1
                    nxt.append(b)
        frontier = nxt
    return dist


def bitmask(indices, n_atoms):
    words = (n_atoms + 63) // 64
    out = np.zeros(words, dtype=np.uint64)
    for i in indices:
        out[i >> 6] |= np.uint64(1) << np.uint64(i & 63)
    return out


@dataclass
class ContactParams:
    """gaudi/objectives/contacts.py:101-121 plus DetectClash's fixed options (:142-145)."""
    radius: float = 5.0
    clash_threshold: float = 0.6
    hydrophobic_threshold: float = -0.4
    hbond_allowance: float = 0.2
    assumed_max_vdw: float = 2.1
    cutoff: float = 0.0
    weight: float = 1.0
    bond_separation: int = 4
    only_internal: bool = False
    same_residue: bool = True   # DetectClash intraRes (contacts.py:142): False drops pairs inside one residue (host prep)
    cull: bool = True       # B200 path only: skip receptor tiles out of reach of a pose
    tiled: bool = False     # B200 path only: two-kernel tiled Contacts path instead of the fused one
    sequential_express: bool = False   # B200 path only: bond rotations one after another (reference order)
    rmsd_simt: bool = False            # B200 path only: similarity bitmask by the FP32 direct-difference kernel
    rmsd_hmma: bool = False            # B200 path only: pre-filter product by mma.sync instead of tcgen05 (cross-check)


@dataclass
class MobileModel:
    xyz0: np.ndarray            # (M, 3) f64 template coordinates (input-file frame)
    radius: np.ndarray          # (M,)  f64
    flags: np.ndarray           # (M,)  u8
    bonds: np.ndarray           # (nb, 2) i32 bonds among mobile atoms
    tors_fixed: np.ndarray      # (T,) i32   anchor-side atom of each rotatable bond
    tors_moving: np.ndarray     # (T,) i32   moving-side atom
    tors_mask: np.ndarray       # (T, W) u64 moving-side atom sets, W = ceil(M / 64)
    tors_snap: np.ndarray       # (T,) u8    both atoms C2/N2 -> 0/180 only (torsion.py:159)
    tors_src: np.ndarray        # (T,) i32   >= 0: column of the torsion allele;
                                #            < 0: rotamer driven, -(1 + 4 * residue + chi)
    excl_mask: np.ndarray       # (M, W) u64 atoms within <= bond_separation bonds (self incl.)
    origin_atom: int = 0        # Search target atom (search.py:186-188)
    torsion_before_search: bool = True
    rot_ptr: np.ndarray = field(default_factory=lambda: np.zeros(1, dtype=np.int32))
    rot_delta: np.ndarray = field(default_factory=lambda: np.zeros((0, 4), dtype=np.float64))

    @property
    def n_atoms(self):
        return int(self.xyz0.shape[0])

    @property
    def n_torsions(self):
        return int(self.tors_fixed.shape[0])

    @property
    def n_allele_torsions(self):
        return int((self.tors_src >= 0).sum())

    @property
    def n_rotamer_residues(self):
        return int(self.rot_ptr.shape[0] - 1)

    def moving_indices(self, t):
        m = self.tors_mask[t]
        return np.array([i for i in range(self.n_atoms)
                         if (int(m[i >> 6]) >> (i & 63)) & 1], dtype=np.int64)


@dataclass
class ReceptorModel:
    xyz: np.ndarray             # (R, 3) f64 world coordinates
    radius: np.ndarray          # (R,)  f64
    flags: np.ndarray           # (R,)  u8 (F_SEED marks CB atoms of Rotamers residues)
    bonds: np.ndarray           # (nb, 2) i32 bonds among static atoms
    rr_ptr: np.ndarray          # (R + 1,) i32 CSR over static atoms a ...
    rr_idx: np.ndarray          # (nnz,)  i32 ... partners b > a, > bond_separation bonds apart
                                #          and close enough to ever be reported by DetectClash
    centre: np.ndarray          # (3,) f64 re-centring origin of the fp32 device frame
    ms_excl_ptr: np.ndarray = field(default_factory=lambda: np.zeros(1, dtype=np.int32))
    ms_excl_idx: np.ndarray = field(default_factory=lambda: np.zeros(0, dtype=np.int32))
    # ^ CSR over *mobile* atoms: static atoms within <= bond_separation bonds (Rotamers only)

    @property
    def n_atoms(self):
        return int(self.xyz.shape[0])


def build_mobile(xyz0, radius, flags, bonds, rotatable, origin_atom=0,
                 torsion_before_search=True, snap=None, bond_separation=4,
                 tors_src=None, rot_ptr=None, rot_delta=None):
    """Flatten a mobile molecule.

    ``rotatable`` is a list of ``(fixed, moving)`` atom-index pairs in the reference's bond
    order (sorted by min serial, torsion.py:246-248); ``fixed`` is the atom on the anchor
    side (``box.find_nearest(anchor, bond.atoms)``, torsion.py:305).
    """
    xyz0 = np.ascontiguousarray(xyz0, dtype=np.float64)
    m = xyz0.shape[0]
    bonds = np.asarray(bonds, dtype=np.int32).reshape(-1, 2)
    adj = adjacency(m, bonds)
    words = (m + 63) // 64
    t = len(rotatable)
    tors_mask = np.zeros((t, words), dtype=np.uint64)
    for k, (fx, mv) in enumerate(rotatable):
        tors_mask[k] = bitmask(side_of_bond(adj, int(fx), int(mv)), m)
    excl = np.zeros((m, words), dtype=np.uint64)
    for a in range(m):
        excl[a] = bitmask(within_bonds(adj, a, bond_separation), m)
    rot = np.asarray(rotatable, dtype=np.int32).reshape(-1, 2)
    return MobileModel(
        xyz0=xyz0,
        radius=np.ascontiguousarray(radius, dtype=np.float64),
        flags=np.ascontiguousarray(flags, dtype=np.uint8),
        bonds=bonds,
        tors_fixed=np.ascontiguousarray(rot[:, 0]),
        tors_moving=np.ascontiguousarray(rot[:, 1]),
        tors_mask=tors_mask,
        tors_snap=(np.zeros(t, dtype=np.uint8) if snap is None
                   else np.ascontiguousarray(snap, dtype=np.uint8)),
        tors_src=(np.arange(t, dtype=np.int32) if tors_src is None
                  else np.ascontiguousarray(tors_src, dtype=np.int32)),
        excl_mask=excl,
        origin_atom=int(origin_atom),
        torsion_before_search=bool(torsion_before_search),
        rot_ptr=(np.zeros(1, dtype=np.int32) if rot_ptr is None
                 else np.ascontiguousarray(rot_ptr, dtype=np.int32)),
        rot_delta=(np.zeros((0, 4), dtype=np.float64) if rot_delta is None
                   else np.ascontiguousarray(rot_delta, dtype=np.float64).reshape(-1, 4)),
    )


def static_pair_candidates(xyz, radius, bonds, params, max_mobile_radius=0.0):
    """CSR of static-static pairs that DetectClash could ever report: more than
    ``bond_separation`` bonds apart and within the loosest geometric bound
    ``d <= r_a + r_b - min(thresholds)``.  Their *values* are computed on the device
    (``gmm_set_receptor``); this is only the neighbour/topology search."""
    from scipy.spatial import cKDTree
    r = xyz.shape[0]
    ptr = np.zeros(r + 1, dtype=np.int32)
    if r < 2:
        return ptr, np.zeros(0, dtype=np.int32)
    thr = min(params.hydrophobic_threshold, params.clash_threshold)
    rmax = float(np.max(radius))
    tree = cKDTree(xyz)
    pairs = tree.query_pairs(2.0 * rmax - thr + 1e-6, output_type='ndarray')
    if len(pairs) == 0:
        return ptr, np.zeros(0, dtype=np.int32)
    a = np.minimum(pairs[:, 0], pairs[:, 1])
    b = np.maximum(pairs[:, 0], pairs[:, 1])
    d = np.sqrt(((xyz[a] - xyz[b]) ** 2).sum(axis=1))
    keep = d <= radius[a] + radius[b] - thr + 1e-6
    a, b = a[keep], b[keep]
    # drop pairs within <= bond_separation bonds
    adj = adjacency(r, bonds)
    cache = {}
    ok = np.ones(len(a), dtype=bool)
    for k in range(len(a)):
        ai = int(a[k])
        if not adj[ai]:
            continue
        nb = cache.get(ai)
        if nb is None:
            nb = cache[ai] = within_bonds(adj, ai, params.bond_separation)
        if int(b[k]) in nb:
            ok[k] = False
    a, b = a[ok], b[ok]
    order = np.lexsort((b, a))
    a, b = a[order], b[order]
    counts = np.bincount(a, minlength=r)
    ptr[1:] = np.cumsum(counts)
    return ptr, b.astype(np.int32)


def build_receptor(xyz, radius, flags, bonds, params, centre=None):
    xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
    radius = np.ascontiguousarray(radius, dtype=np.float64)
    bonds = np.asarray(bonds, dtype=np.int32).reshape(-1, 2)
    if params.only_internal:
        ptr, idx = np.zeros(xyz.shape[0] + 1, dtype=np.int32), np.zeros(0, dtype=np.int32)
    else:
        ptr, idx = static_pair_candidates(xyz, radius, bonds, params)
    if centre is None:
        centre = xyz.mean(axis=0) if len(xyz) else np.zeros(3)
    return ReceptorModel(xyz=xyz, radius=radius,
                         flags=np.ascontiguousarray(flags, dtype=np.uint8),
                         bonds=bonds, rr_ptr=ptr, rr_idx=idx,
                         centre=np.ascontiguousarray(centre, dtype=np.float64))


def split_molecule(xyz, radius, flags, bonds, mobile_idx, params, rotatable=(), origin_atom=None,
                   centre=None, **mobile_kw):
    """Split one bonded system into a mobile part (atoms ``mobile_idx``, in that order) and
    a static part (everything else), keeping DetectClash's bond-separation exclusions across
    the cut (``ms_excl``).  This is how flexible side chains of a Rotamers gene sit next to
    the rest of their protein.  ``rotatable`` uses indices of the *whole* system.
    Returns ``(MobileModel, ReceptorModel, static_idx)``."""
    xyz = np.asarray(xyz, dtype=np.float64)
    n = xyz.shape[0]
    bonds = np.asarray(bonds, dtype=np.int64).reshape(-1, 2)
    mobile_idx = np.asarray(mobile_idx, dtype=np.int64)
    is_mobile = np.zeros(n, dtype=bool)
    is_mobile[mobile_idx] = True
    static_idx = np.flatnonzero(~is_mobile)
    to_m = -np.ones(n, dtype=np.int64)
    to_m[mobile_idx] = np.arange(len(mobile_idx))
    to_s = -np.ones(n, dtype=np.int64)
    to_s[static_idx] = np.arange(len(static_idx))
    mm = bonds[is_mobile[bonds[:, 0]] & is_mobile[bonds[:, 1]]]
    ss = bonds[~is_mobile[bonds[:, 0]] & ~is_mobile[bonds[:, 1]]]
    cross = bonds[is_mobile[bonds[:, 0]] != is_mobile[bonds[:, 1]]]
    # rotatable bonds may have their fixed atom in the static part: the moving side is then
    # computed on the full graph and must lie entirely in the mobile part
    adj = adjacency(n, bonds)
    m = len(mobile_idx)
    words = (m + 63) // 64
    mob = build_mobile(xyz[mobile_idx], np.asarray(radius)[mobile_idx], np.asarray(flags)[mobile_idx],
                       np.stack([to_m[mm[:, 0]], to_m[mm[:, 1]]], axis=1) if len(mm) else np.zeros((0, 2), np.int32),
                       [], origin_atom=0 if origin_atom is None else int(to_m[origin_atom]),
                       bond_separation=params.bond_separation, **mobile_kw)
    if len(rotatable):
        fixed, moving, masks = [], [], []
        for fx, mv in rotatable:
            side = side_of_bond(adj, int(fx), int(mv))
            if not all(is_mobile[a] for a in side):
                raise ValueError("moving side of bond %d-%d leaves the mobile part" % (fx, mv))
            if to_m[fx] < 0:
                raise ValueError("fixed atom %d of a rotatable bond must be in the mobile part "
                                 "(add it to mobile_idx; it never moves)" % fx)
            fixed.append(int(to_m[fx])); moving.append(int(to_m[mv]))
            masks.append(bitmask([int(to_m[a]) for a in side], m))
        mob.tors_fixed = np.asarray(fixed, np.int32)
        mob.tors_moving = np.asarray(moving, np.int32)
        mob.tors_mask = np.asarray(masks, np.uint64).reshape(len(fixed), words)
        mob.tors_snap = np.zeros(len(fixed), np.uint8)
        if "tors_src" not in mobile_kw or mobile_kw["tors_src"] is None:
            mob.tors_src = np.arange(len(fixed), dtype=np.int32)
    # exclusions across the cut, on the full graph
    excl = np.zeros((m, words), dtype=np.uint64)
    ms_ptr = np.zeros(m + 1, dtype=np.int32)
    ms_idx = []
    for k, a in enumerate(mobile_idx):
        near = within_bonds(adj, int(a), params.bond_separation)
        excl[k] = bitmask([int(to_m[b]) for b in near if is_mobile[b]], m)
        ms_idx.extend(sorted(int(to_s[b]) for b in near if not is_mobile[b]))
        ms_ptr[k + 1] = len(ms_idx)
    mob.excl_mask = excl
    rec = build_receptor(xyz[static_idx], np.asarray(radius)[static_idx], np.asarray(flags)[static_idx],
                         np.stack([to_s[ss[:, 0]], to_s[ss[:, 1]]], axis=1) if len(ss) else np.zeros((0, 2), np.int32),
                         params, centre=xyz.mean(axis=0) if centre is None else centre)
    # static-static exclusions must also see paths that run through mobile atoms
    if len(cross) and len(rec.rr_idx):
        keep = np.ones(len(rec.rr_idx), dtype=bool)
        touched = {int(to_s[b]) for a in mobile_idx for b in within_bonds(adj, int(a), params.bond_separation)
                   if not is_mobile[b]}
        for sa in touched:
            near = within_bonds(adj, int(static_idx[sa]), params.bond_separation)
            for e in range(rec.rr_ptr[sa], rec.rr_ptr[sa + 1]):
                if int(static_idx[rec.rr_idx[e]]) in near:
                    keep[e] = False
        for sb in touched:          # rows a < b where b is the touched atom
            near = within_bonds(adj, int(static_idx[sb]), params.bond_separation)
            for sa in {int(to_s[x]) for x in near if not is_mobile[x]}:
                for e in range(rec.rr_ptr[sa], rec.rr_ptr[sa + 1]):
                    if rec.rr_idx[e] == sb:
                        keep[e] = False
        counts = np.array([keep[rec.rr_ptr[a]:rec.rr_ptr[a + 1]].sum() for a in range(rec.n_atoms)])
        rec.rr_idx = rec.rr_idx[keep]
        rec.rr_ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    rec.ms_excl_ptr = ms_ptr
    rec.ms_excl_idx = np.asarray(ms_idx, dtype=np.int32)
    rec.ms_bonds = np.array([(to_m[a], to_s[b]) if is_mobile[a] else (to_m[b], to_s[a]) for a, b in cross],
                            dtype=np.int32).reshape(-1, 2)
    return mob, rec, static_idx


def empty_receptor(centre=(0.0, 0.0, 0.0)):
    return ReceptorModel(xyz=np.zeros((0, 3)), radius=np.zeros(0), flags=np.zeros(0, np.uint8),
                         bonds=np.zeros((0, 2), np.int32), rr_ptr=np.zeros(1, np.int32),
                         rr_idx=np.zeros(0, np.int32), centre=np.asarray(centre, np.float64))


def exclude_same_residue(mob, rec, mobile_residue, static_residue):
    """``Contacts(same_residue=False)`` = DetectClash ``intraRes=False`` (gaudi/objectives/contacts.py:142):
    pairs of atoms of one residue are never reported.  Done entirely in the static model, in place:
    mobile-mobile pairs through ``excl_mask``, static-static pairs dropped from the candidate CSR,
    mobile-static pairs (a flexible side chain against its own backbone) added to ``ms_excl``.
    ``mobile_residue`` / ``static_residue``: one integer per atom, unique per (molecule, residue)."""
    mres = np.asarray(mobile_residue, dtype=np.int64)
    sres = np.asarray(static_residue, dtype=np.int64)
    m = mob.n_atoms
    for a in range(m):
        same = np.flatnonzero(mres == mres[a])
        for b in same:
            mob.excl_mask[a, int(b) >> 6] |= np.uint64(1) << np.uint64(int(b) & 63)
    if len(rec.rr_idx):
        rows = np.repeat(np.arange(rec.n_atoms), np.diff(rec.rr_ptr))
        keep = sres[rows] != sres[rec.rr_idx]
        counts = np.bincount(rows[keep], minlength=rec.n_atoms)
        rec.rr_idx = rec.rr_idx[keep]
        rec.rr_ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    by_res = {}
    for j, r in enumerate(sres):
        by_res.setdefault(int(r), []).append(j)
    old_ptr = rec.ms_excl_ptr if len(rec.ms_excl_ptr) == m + 1 else np.zeros(m + 1, dtype=np.int32)
    ptr, idx = np.zeros(m + 1, dtype=np.int32), []
    for a in range(m):
        have = set(int(v) for v in rec.ms_excl_idx[old_ptr[a]:old_ptr[a + 1]])
        have.update(by_res.get(int(mres[a]), ()))
        idx.extend(sorted(have))
        ptr[a + 1] = len(idx)
    rec.ms_excl_ptr, rec.ms_excl_idx = ptr, np.asarray(idx, dtype=np.int32)
    mob.residue, rec.residue = mres, sres          # (read by the tests' oracle adapter)
    return mob, rec
