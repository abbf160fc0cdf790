"""
CPU oracle for the GaudiMM hot path (TEST INFRASTRUCTURE — not product code).

This module is a numpy / fp64 restatement of the reference's CPU path for

    genome expression (Torsion, Rotamers, Search)  ->  Contacts objective  ->  similarity.rmsd

It exists only so that ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` can check and time the CUDA
path against the reference's algorithm.  Nothing under ``gaudi_b200/`` may import it.

Parity status: PINNED.  ``tests/test_oracle_golden.py`` checks this file against every
hot-path golden value of the reference's own test-suite
through fixtures generated from ``/root/reference/tests/data`` by
``tests/golden/make_golden.py``; ``rmsd_squared`` is additionally checked against the
reference's own ``gaudi/similarity.py`` imported in the build container.

The geometry primitives of the reference live in UCSF Chimera (not vendored, not
installable here): ``BondRot.adjustAngle``, ``openState.xform``, ``specifier.zone`` and
``DetectClash.detectClash``.  Their published behaviour is restated below; each function
cites the reference call site (paths relative to the reference repository root) that
fixes its semantics.
"""
from __future__ import annotations

import math

import numpy as np

# Flag bits shared with the product's static model (gaudi_b200/model.py).  Re-declared
# here on purpose: the oracle must not import the product.
F_HYDROPHOBIC = 1   # element in Contacts.hydrophobic_elements (contacts.py:103)
F_DONOR = 2         # H-bond donor (DetectClash hbondAllowance rule)
F_ACCEPTOR = 4      # H-bond acceptor
F_SEED = 8          # zone seed: probe atom or CB of a Rotamers residue (contacts.py:195-207)
F_XFORM = 16        # atom belongs to the molecule the Search gene transforms

PI_LITERAL = 3.14159265359  # contacts.py:255 uses this literal, not math.pi


# --------------------------------------------------------------------------------------
# Genome expression
# --------------------------------------------------------------------------------------
def rotate_about_bond(coords, fixed, moving, moving_idx, angle_deg):
    """In-place right-handed rotation of ``coords[moving_idx]`` by ``angle_deg`` about the
    axis pointing from atom ``fixed`` to atom ``moving``.

    Restates ``chimera.BondRot.adjustAngle(delta, anchor)`` as used at
    gaudi/genes/torsion.py:161 and gaudi/genes/rotamers.py:189: the side of the bond that
    holds the anchor stays put, the other side turns.  The sign convention is pinned by the
    golden ``58.986905376603815`` (tests/test_genes_torsion.py:54-63).
    """
    if angle_deg == 0.0 or len(moving_idx) == 0:
        return
    a = coords[fixed]
    b = coords[moving]
    u = b - a
    u = u / np.sqrt(np.dot(u, u))
    th = math.radians(angle_deg)
    c, s = math.cos(th), math.sin(th)
    v = coords[moving_idx] - b
    cross = np.cross(np.broadcast_to(u, v.shape), v)
    dot = v @ u
    coords[moving_idx] = b + v * c + cross * s + np.outer(dot, u) * (1.0 - c)


def torsion_angle(alpha, snap):
    """gaudi/genes/torsion.py:159-160 — bonds whose two atoms are both C2/N2 only take the
    values 0 or 180 degrees."""
    if snap:
        return 0.0 if alpha <= 0 else 180.0
    return float(alpha)


def express(xyz0, torsions, torsion_allele, xform=None, origin_atom=0,
            torsion_before_search=True, xform_mask=None,
            rotamer_torsions=None, rotamer_allele=None, rotamer_tables=None):
    """Expressed coordinates of one individual (fp64, world frame).

    Parameters
    ----------
    xyz0 : (M, 3) template (input-file) coordinates of the mobile atoms.
    torsions : list of (fixed, moving, moving_idx, snap) in the reference's bond order
        (gaudi/genes/torsion.py:245-248).
    torsion_allele : (T,) degrees.  They are *deltas from the input geometry*: the
        reference calls ``br.adjustAngle(alpha - br.angle, ...)`` with ``br.angle == 0``
        at load time (torsion.py:157-161).
    xform : (3, 4) or (12,) product ``T . R`` of the Search allele (search.py:201-214).
        ``Search.express`` sets ``xform = T . R . to_zero(origin)``, i.e.
        ``x' = R (x - origin) + t``.
    origin_atom : index of the Search target atom; its *current local* coordinate is the
        origin (search.py:186-199), so it is taken after the torsions when Torsion is
        listed before Search in the YAML (base.py:157-168 expresses genes in order).
    xform_mask : (M,) bool — atoms of the molecule Search acts on (others keep identity).
    rotamer_torsions : list of (fixed, moving, moving_idx, residue, chi_index);
    rotamer_allele : (n_res,) floats in [0, 1); rotamer_tables : list of (n_rot, 4) arrays
        of chi *deltas* (chi - chi0) in degrees: ``rotamers[int(i * len(rotamers))]``
        (rotamers.py:141-152), ``adjustAngle(chi - bondrot.chi)`` (rotamers.py:186-189).
    """
    x = np.array(xyz0, dtype=np.float64, copy=True)
    origin_pre = x[origin_atom].copy()
    # Rotamers come before Torsion/Search in every reference example; they act on other
    # atoms than the ligand torsions, so the relative order does not change the result.
    if rotamer_torsions:
        for fixed, moving, idx, res, chi in rotamer_torsions:
            table = rotamer_tables[res]
            k = int(rotamer_allele[res] * len(table))
            rotate_about_bond(x, fixed, moving, idx, float(table[k][chi]))
    for (fixed, moving, idx, snap), alpha in zip(torsions, torsion_allele):
        rotate_about_bond(x, fixed, moving, idx, torsion_angle(alpha, snap))
    if xform is not None:
        m = np.asarray(xform, dtype=np.float64).reshape(3, 4)
        origin = x[origin_atom].copy() if torsion_before_search else origin_pre
        sel = slice(None) if xform_mask is None else np.asarray(xform_mask, dtype=bool)
        x[sel] = (x[sel] - origin) @ m[:, :3].T + m[:, 3]
    return x


def multiply_3x4(*mats):
    """``Matrix.multiply_matrices`` of Chimera as used at search.py:206-214: affine product
    of 3x4 matrices, right-most applied first."""
    out = np.eye(4)
    for m in mats:
        a = np.eye(4)
        a[:3, :] = np.asarray(m, dtype=np.float64).reshape(3, 4)
        out = out @ a
    return out[:3, :]


def dihedral(p0, p1, p2, p3):
    """IUPAC dihedral in degrees (``chimera.dihedral``), used for chi angles at
    rotamers.py:203; pinned by tests/test_genes_rotamers.py:44-45."""
    b0 = p0 - p1
    b1 = p2 - p1
    b2 = p3 - p2
    b1n = b1 / np.linalg.norm(b1)
    v = b0 - np.dot(b0, b1n) * b1n
    w = b2 - np.dot(b2, b1n) * b1n
    x = np.dot(v, w)
    y = np.dot(np.cross(b1n, v), w)
    return math.degrees(math.atan2(y, x))


# --------------------------------------------------------------------------------------
# Contacts objective
# --------------------------------------------------------------------------------------
def lennard_jones(r1, r2, overlap):
    """gaudi/objectives/contacts.py:215-241 (distance rebuilt from the overlap, so it carries
    the H-bond allowance when one was applied)."""
    r0 = r1 + r2
    distance = r0 - overlap
    x = (r0 / distance) ** 6
    return x * x - 2 * x


def vdw_vol_overlap(r1, r2, overlap):
    """gaudi/objectives/contacts.py:243-268."""
    d = r1 + r2 - overlap
    if d == 0:
        return 1000
    h_a, h_b = 0, 0
    if d < (r1 + r2):
        h_a = (r2 ** 2 - (d - r1) ** 2) / (2 * d)
        h_b = (r1 ** 2 - (d - r2) ** 2) / (2 * d)
    return (PI_LITERAL / 3) * ((h_a ** 2) * (3 * r1 - h_a) + (h_b ** 2) * (3 * r2 - h_b))


class Graph(object):
    """Bond graph with cached "within n bonds" neighbourhoods (DetectClash's
    ``bondSeparation`` exclusion, contacts.py:142-145)."""

    def __init__(self, n_atoms, bonds):
        self.adj = [[] for _ in range(n_atoms)]
        for a, b in bonds:
            a, b = int(a), int(b)
            self.adj[a].append(b)
            self.adj[b].append(a)
        self._cache = {}

    def within(self, src, depth):
        key = (src, depth)
        hit = self._cache.get(key)
        if hit is not None:
            return hit
        seen = {src}
        frontier = [src]
        for _ in range(depth):
            nxt = []
            for a in frontier:
                for b in self.adj[a]:
                    if b not in seen:
                        seen.add(b)
                        nxt.append(b)
            frontier = nxt
            if not frontier:
                break
        self._cache[key] = seen
        return seen


def zone_atoms(coords, flags, radius, only_internal=False):
    """Atoms inside the search zone (contacts.py:190-213 + ``chimera.specifier.zone``):
    the seeds themselves plus every atom within ``radius`` of a seed.  With
    ``only_internal`` the zone is just the seeds (contacts.py:199)."""
    seeds = np.flatnonzero(flags & F_SEED)
    if only_internal or len(seeds) == 0:
        return seeds
    from scipy.spatial import cKDTree
    tree = cKDTree(coords)
    near = tree.query_ball_point(coords[seeds], radius)
    inz = np.zeros(len(coords), dtype=bool)
    inz[seeds] = True
    for lst in near:
        inz[lst] = True
    # cKDTree's ball test is d <= r, matching find_close_points.
    return np.flatnonzero(inz)


def detect_clash(coords, radius, flags, atoms, graph, threshold, hbond_allowance=0.2,
                 assumed_max_vdw=2.1, bond_separation=4, residue=None):
    """``DetectClash.detectClash(atoms, test=atoms, clashThreshold=threshold,
    hbondAllowance=0.2, assumedMaxVdw=2.1, bondSeparation=4, intraRes=True,
    interSubmodel=True)`` (contacts.py:140-146).

    Returns ``{(a, b): overlap}`` for a < b; the reference's dict-of-dicts holds every
    such pair twice (once per direction)."""
    from scipy.spatial import cKDTree
    atoms = np.asarray(atoms, dtype=np.int64)
    if len(atoms) < 2:
        return {}
    sub = coords[atoms]
    tree = cKDTree(sub)
    rmax = float(radius[atoms].max())
    # Every pair that can be found from either side lies within this bound.
    pairs = tree.query_pairs(rmax + assumed_max_vdw - threshold, output_type='ndarray')
    out = {}
    for ia, ib in pairs:
        a, b = int(atoms[ia]), int(atoms[ib])
        if a > b:
            a, b = b, a
        d = float(np.sqrt(((coords[a] - coords[b]) ** 2).sum()))
        ra, rb = float(radius[a]), float(radius[b])
        # found from a if d <= ra + assumedMaxVdw - thr, from b likewise
        if d > max(ra, rb) + assumed_max_vdw - threshold:
            continue
        if b in graph.within(a, bond_separation):
            continue
        if residue is not None and residue[a] == residue[b]:      # intraRes=False (contacts.py:142, same_residue)
            continue
        clash = ra + rb - d
        if hbond_allowance:
            fa, fb = int(flags[a]), int(flags[b])
            if ((fa & F_DONOR) and (fb & F_ACCEPTOR)) or ((fb & F_DONOR) and (fa & F_ACCEPTOR)):
                clash -= hbond_allowance
        if clash < threshold:
            continue
        out[(a, b)] = clash
    return out


def analyze_interactions(clashes, radius, flags, clash_threshold):
    """contacts.py:148-188 on unique pairs; the caller doubles the sums because the
    reference iterates both (a1, a2) and (a2, a1)."""
    positive, negative = [], []
    for (a, b), overlap in clashes.items():
        if overlap <= clash_threshold:
            if (flags[a] & F_HYDROPHOBIC) and (flags[b] & F_HYDROPHOBIC):
                positive.append((a, b, overlap, lennard_jones(float(radius[a]), float(radius[b]), overlap)))
        else:
            negative.append((a, b, overlap, vdw_vol_overlap(float(radius[a]), float(radius[b]), overlap)))
    return positive, negative


def contacts(coords, radius, flags, graph, radius_zone=5.0, clash_threshold=0.6,
             hydrophobic_threshold=-0.4, cutoff=0.0, weight=1.0, bond_separation=4,
             only_internal=False, hbond_allowance=0.2, assumed_max_vdw=2.1,
             return_pairs=False, residue=None):
    """Both Contacts scores of one expressed individual (``residue``: per-atom residue ids when
    ``same_residue`` is False, contacts.py:142 ``intraRes``).

    ``coords/radius/flags`` cover every atom of every Molecule gene (mobile first, then
    static, by the convention of the tests).  Returns ``(clashes, hydrophobic)`` exactly
    as ``evaluate_clashes`` / ``evaluate_hydrophobic`` would (contacts.py:129-138), before
    the caller's ``round(score, precision)`` (base.py:361)."""
    coords = np.asarray(coords, dtype=np.float64)
    atoms = zone_atoms(coords, flags, radius_zone, only_internal)
    # which == 'clashes': threshold = clash_threshold (contacts.py:120-121)
    found_c = detect_clash(coords, radius, flags, atoms, graph, clash_threshold,
                           hbond_allowance, assumed_max_vdw, bond_separation, residue)
    _, neg = analyze_interactions(found_c, radius, flags, clash_threshold)
    clashscore = 2.0 * sum(abs(v) for (_, _, _, v) in neg)
    if cutoff and clashscore > cutoff:
        clashscore = -1000 * weight
    # which == 'hydrophobic': threshold = hydrophobic_threshold (contacts.py:116-118)
    found_h = detect_clash(coords, radius, flags, atoms, graph, hydrophobic_threshold,
                           hbond_allowance, assumed_max_vdw, bond_separation, residue)
    pos, _ = analyze_interactions(found_h, radius, flags, clash_threshold)
    hydro = 2.0 * sum(v for (_, _, _, v) in pos)
    if return_pairs:
        return clashscore, hydro, neg, pos
    return clashscore, hydro


# --------------------------------------------------------------------------------------
# similarity.rmsd
# --------------------------------------------------------------------------------------
def rmsd_squared(coords1, coords2):
    """gaudi/similarity.py:73-75 — no superposition."""
    diff = coords1 - coords2
    return (diff * diff).sum() / coords1.shape[0]


def rmsd_similar(coords1, coords2, threshold):
    """gaudi/similarity.py:62-70 for one subject molecule."""
    if coords1.shape[0] != coords2.shape[0]:
        return False
    return not (rmsd_squared(coords1, coords2) > threshold * threshold)


def rmsd_matrix(xyz, rows=None):
    """All-pairs RMSD of ``xyz`` (N, A, 3) restricted to ``rows`` (vectorised form of
    calling ``rmsd_squared`` per pair as gaudi/algorithms.py:141-155 does)."""
    xyz = np.asarray(xyz, dtype=np.float64)
    sel = xyz if rows is None else xyz[rows]
    out = np.empty((sel.shape[0], xyz.shape[0]))
    for i in range(sel.shape[0]):
        d = xyz - sel[i]
        out[i] = np.sqrt((d * d).sum(axis=(1, 2)) / xyz.shape[1])
    return out


# --------------------------------------------------------------------------------------
# Greedy niching pass (gaudi/algorithms.py:139-155), used to check the host-side pass that
# reads the device similarity bitmask.
# --------------------------------------------------------------------------------------
def greedy_niching(order, similar, not_better):
    """``order``: candidate ids in ``population + offspring`` order; ``similar(i, j)`` and
    ``not_better(i, j)`` (``i.fitness <= j.fitness``) are callables.  Returns kept ids."""
    kept = []
    for i in order:
        to_delete = []
        broke = False
        for pos, n_i in enumerate(kept):
            if similar(i, n_i):
                if not_better(i, n_i):
                    broke = True
                    break
                to_delete.append(pos)
        if not broke:
            if to_delete:
                kept = [k for j, k in enumerate(kept) if j not in to_delete]
            kept.append(i)
    return kept


# --------------------------------------------------------------------------------------
# Geometric restraint objectives ("next" row of the scope)
# Parity status of this section: UNPINNED.  The reference has no test for its Distance and Angle
# objectives, so there is no golden value to reproduce; the two functions below follow the
# reference's own lines (cited) and are checked on hand-computable geometry only.
# --------------------------------------------------------------------------------------
def distance_objective(probe_xyz, target_xyz, threshold, tolerance=None, weight=1.0):
    """gaudi/objectives/distance.py:102-129 (``evaluate_distances``): mean absolute deviation of
    the probe-target distances from ``threshold``; a probe closer than ``threshold + tolerance``
    contributes ``-1000 * weight`` before the absolute value."""
    distances = []
    target = np.asarray(target_xyz, dtype=np.float64)
    for a in np.asarray(probe_xyz, dtype=np.float64).reshape(-1, 3):
        d = math.sqrt(((a - target) ** 2).sum()) - threshold
        if tolerance is not None and d < tolerance:
            distances.append(-1000 * weight)
        else:
            distances.append(d)
    return float(np.mean(np.absolute(distances)))


def distance_center_of_mass(probe_xyz, masses, target_xyz):
    """gaudi/objectives/distance.py:131-148 (``evaluate_center_of_mass``): distance between the target
    and the mass-weighted centre of the probe atoms (no threshold, no tolerance)."""
    centre = np.average(np.asarray(probe_xyz, dtype=np.float64).reshape(-1, 3), axis=0, weights=masses)
    return float(np.sqrt(((centre - np.asarray(target_xyz, dtype=np.float64)) ** 2).sum()))


def angle_objective(points, threshold):
    """gaudi/objectives/angle.py:86-96: three points -> ``chimera.angle`` (angle at the middle
    point), four -> ``chimera.dihedral``; ``'planar'`` -> |sin(angle)|, else |threshold - angle|."""
    p = [np.asarray(x, dtype=np.float64) for x in points]
    if len(p) == 3:
        u, w = p[0] - p[1], p[2] - p[1]
        c = np.dot(u, w) / (np.linalg.norm(u) * np.linalg.norm(w))
        angle = math.degrees(math.acos(max(-1.0, min(1.0, c))))
    else:
        angle = dihedral(*p)
    if threshold == 'planar':
        return abs(math.sin(math.radians(angle)))
    return abs(threshold - angle)
