"""
CPU oracle for the Coordination objective (TEST INFRASTRUCTURE — not product code).

numpy / fp64 restatement of ``gaudi/objectives/coordination.py`` (sphere search :233-283,
evaluate :180-231, ``ideal_bond_deviation`` :286-360, ``ideal_bonded_positions`` :363-372) and of
the rigid branch of ``gaudi/_cpdrift.py`` (:45-93, :127-199) that it calls.  Only ``tests/`` and
``__graft_entry__.smoke()`` may import it; nothing under ``gaudi_b200/`` does.

Parity status: PINNED.
  * both golden scores of the reference's own test (tests/test_objectives_coordination.py:33-60,
    4c3w.mol2, metal 1005, octahedron: 2.4696 and, with both corrections, 3.2405) and the ligand
    set it asserts (serials 1013, 1014, 1021, 1023, 1024, 408) are reproduced from the fixture
    ``tests/golden/4c3w_mol2.npz`` (tests/test_oracle_golden.py);
  * ``rigid_cpd_rmsd`` is checked against the reference's own ``gaudi/_cpdrift.py``, which runs
    under Python 3 in the build container: ``tests/golden/cpd_reference.npz`` holds its outputs on
    seeded point sets (generator: tests/golden/make_golden.py).

Three facts of UCSF Chimera (not vendored) enter the score; each is restated here and fixed by the
goldens above:
  * ``chimera.Element.bondLength(a, b)`` = sum of the two covalent bond radii (V 1.33, O 0.68,
    N 0.68 give the 0.5964 the corrected golden needs for its distance term);
  * ``chimera.idatm.typeInfo[type].geometry``: 4 tetrahedral, 3 planar, 2 linear, 1 single;
  * ``chimera.bondGeom.bondPositions`` (free valence positions of an atom).  The golden fixes the
    planar case with one substituent: positions at 120 degrees from the bond, in the plane spanned
    by the bond and the y axis, the +y side first.  The tetrahedral case with two substituents has
    two mirror-image positions and nothing in the reference's tests says which comes first: the
    choice below is a convention (it affects the score only through the dihedral term of such
    ligand atoms, e.g. a thioether sulphur).
"""
from __future__ import annotations

import math

import numpy as np

GEOMETRIES = {
    # coordination.py:51-64; the last point is the metal
    "tetrahedral": [(1.0, 1.0, 1.0), (-1.0, -1.0, 1.0), (-1.0, 1.0, -1.0), (1.0, -1.0, -1.0), (0.0, 0.0, 0.0)],
    "octahedron": [(1.0, 0.0, 0.0), (-1.0, 0.0, 0.0), (0.0, 1.0, 0.0), (0.0, -1.0, 0.0), (0.0, 0.0, 1.0),
                   (0.0, 0.0, -1.0), (0.0, 0.0, 0.0)],
    "square planar": [(1.0, 0.0, 0.0), (-1.0, 0.0, 0.0), (0.0, 1.0, 0.0), (0.0, -1.0, 0.0), (0.0, 0.0, 0.0)],
    "linear": [(0.0, 1.0, 0.0), (0.0, -1.0, 0.0), (0.0, 0.0, 0.0)],
    "trigonal planar": [(1.0, 0.0, 0.0), (-0.5, 0.0, 0.8660), (-0.5, 0.0, -0.8660), (0.0, 0.0, 0.0)],
}


# --------------------------------------------------------------------------------------
# rigid coherent point drift (gaudi/_cpdrift.py)
# --------------------------------------------------------------------------------------
def start_rotations(steps):
    """_cpdrift.py:168-185: ``steps``^3 rotations spread over SO(3) (Shoemake's uniform
    quaternions on a regular grid), each turned into (angle, axis) and then into a matrix."""
    out = []
    for ix in range(steps):
        x = ix / steps
        for i0 in range(steps):
            t0 = 2.0 * math.pi * i0 / steps
            for i1 in range(steps):
                t1 = 2.0 * math.pi * i1 / steps
                r0, r1 = math.sqrt(1.0 - x), math.sqrt(x)
                qs, qi, qj, qk = math.sin(t0) * r0, math.cos(t0) * r0, math.sin(t1) * r1, math.cos(t1) * r1
                theta = 2.0 * math.acos(qs)                      # Quaternion.axis_angle, :334-342
                try:
                    ax = tuple(v / math.sin(theta / 2.0) for v in (qi, qj, qk))
                except ZeroDivisionError:
                    ax = (1.0, 0.0, 0.0)
                c, s = math.cos(theta), math.sin(theta)
                ux, uy, uz = ax
                out.append(np.array([[c + ux * ux * (1 - c), ux * uy * (1 - c) - uz * s, (1 - c) * ux * uz + uy * s],
                                     [uy * ux * (1 - c) + uz * s, c + uy * uy * (1 - c), uy * uz * (1 - c) - ux * s],
                                     [uz * ux * (1 - c) - uy * s, uz * uy * (1 - c) + ux * s, c + uz * uz * (1 - c)]]))
    return out


def _spread(p):
    return math.sqrt(np.var(p, axis=0).mean())            # _cpdrift.py:156-157


def _sqdist(a, b):
    return ((a[:, None, :] - b[None, :, :]) ** 2).sum(axis=2)


def rigid_cpd_rmsd(X, Y, guess_steps=2, max_iterations=10):
    """``coherent_point_drift(X, Y, method='rigid', guess_steps, max_iterations)[2]``
    (_cpdrift.py:45-60) with w = 0: for every start rotation up to ``max_iterations`` EM steps of
    rigid CPD (:63-93, :127-149); a step whose arithmetic turns invalid (0/0, a negative variance
    under the power, every overlap underflowed) ends that start, as numpy's ``invalid='raise'``
    does there.  Returns ``(rmsd, ok)``; ``ok`` is False when some start did not complete a
    single step — the reference then fails inside ``last()`` and Coordination.evaluate returns
    its -1000 penalty (coordination.py:206-209)."""
    X = np.asarray(X, np.float64)
    Y = np.asarray(Y, np.float64)
    N, M_, D = len(X), len(Y), 3
    best = 1000.0
    sx = _spread(X)
    for R0 in start_rotations(guess_steps):
        R = R0
        with np.errstate(all="ignore"):
            s = sx / _spread(Y.dot(R.T))
            t = X.mean(axis=0) - (s * Y.dot(R.T)).mean(axis=0)
            sigma2 = _sqdist(s * Y.dot(R.T) + t, X).sum() / (D * M_ * N)
            done = 0
            for _ in range(max_iterations):
                TY = s * Y.dot(R.T) + t
                if not (sigma2 > 0.0):       # negative: (2 pi sigma2) ** 1.5 is invalid; zero: 0 / 0 or 0 / (0 + 0)
                    break
                dist = _sqdist(TY, X)
                overlap = np.exp(-dist / (2.0 * sigma2))
                col = overlap.sum(axis=0)
                if (col == 0.0).any() or not np.isfinite(col).all():     # 0 / (0 + 0)
                    break
                P = overlap / col
                Np = P.sum()
                mu_x = X.T.dot(P.sum(axis=0)) / Np
                mu_y = Y.T.dot(P.sum(axis=1)) / Np
                Xh, Yh = X - mu_x, Y - mu_y
                A = Xh.T.dot(P.T).dot(Yh)
                U, _, VT = np.linalg.svd(A)
                C = np.eye(D)
                C[-1, -1] = np.linalg.det(U.dot(VT))
                R = U.dot(C).dot(VT)
                tr_ar = np.trace(A.T.dot(R))
                s = tr_ar / np.trace(Yh.T.dot(np.diag(P.sum(axis=1))).dot(Yh))
                t = mu_x - s * R.dot(mu_y)
                sigma2 = (np.trace(Xh.T.dot(np.diag(P.sum(axis=0))).dot(Xh)) - s * tr_ar) / (Np * D)
                if not (np.isfinite(s) and np.isfinite(t).all()):
                    return 1000.0, False     # raised outside the try block of :79-86: evaluate's penalty
                done += 1
            if done == 0:
                return 1000.0, False
            TY = s * Y.dot(R.T) + t
            rmsd = math.sqrt(_sqdist(X, TY).min(axis=1).mean()) / sx            # _cpdrift.py:193-198
        if rmsd < best:
            best = rmsd
    return best, True


# --------------------------------------------------------------------------------------
# chimera.bondGeom.bondPositions, chimera.angle, chimera.dihedral
# --------------------------------------------------------------------------------------
def _unit(v):
    return v / math.sqrt(float(np.dot(v, v)))


def _perpendicular(x):
    """The direction perpendicular to ``x`` in the plane of ``x`` and the y axis (the x axis when
    ``x`` is along y): pinned by the golden 2.4696 for planar atoms with one substituent."""
    for e in (np.array([0.0, 1.0, 0.0]), np.array([1.0, 0.0, 0.0])):
        y = e - np.dot(e, x) * x
        if np.dot(y, y) > 1e-12:
            return _unit(y)
    raise AssertionError


def bond_positions(centre, geometry, length, bonded):
    """Free valence positions of an atom at ``centre`` with substituents at ``bonded``
    (coordination.py:363-372 -> ``chimera.bondGeom.bondPositions``)."""
    nb = len(bonded)
    if geometry < 1 or nb >= geometry:
        return []
    u = [_unit(np.asarray(b, np.float64) - centre) for b in bonded]
    if nb == 0:
        return [centre + length * np.array([1.0, 0.0, 0.0])]          # only its being non-empty is used
    x = -u[0]
    if geometry == 2:
        return [centre + length * x]
    if geometry == 3:
        if nb == 2:
            return [centre - length * _unit(u[0] + u[1])]
        y = _perpendicular(x)
        c, s = math.cos(math.radians(60.0)), math.sin(math.radians(60.0))
        return [centre + length * (c * x + s * y), centre + length * (c * x - s * y)]
    # tetrahedral
    if nb == 3:
        return [centre - length * _unit(u[0] + u[1] + u[2])]
    if nb == 2:
        v = -_unit(u[0] + u[1])
        w = _unit(np.cross(u[0], u[1]))
        half = math.radians(109.4712206 / 2.0)
        return [centre + length * (math.cos(half) * v + math.sin(half) * w),
                centre + length * (math.cos(half) * v - math.sin(half) * w)]
    y = _perpendicular(x)
    a = math.radians(180.0 - 109.4712206)
    pos = []
    for k in range(3):
        phi = 2.0 * math.pi * k / 3.0
        z = np.cross(x, y)
        pos.append(centre + length * (math.cos(a) * x + math.sin(a) * (math.cos(phi) * y + math.sin(phi) * z)))
    return pos


def angle(a, b, c):
    u, v = a - b, c - b
    cs = float(np.dot(u, v)) / math.sqrt(float(np.dot(u, u)) * float(np.dot(v, v)))
    return math.degrees(math.acos(max(-1.0, min(1.0, cs))))


def dihedral(p0, p1, p2, p3):
    b0, b1, b2 = p0 - p1, p2 - p1, p3 - p2
    b1n = _unit(b1)
    v = b0 - np.dot(b0, b1n) * b1n
    w = b2 - np.dot(b2, b1n) * b1n
    return math.degrees(math.atan2(float(np.dot(np.cross(b1n, v), w)), float(np.dot(v, w))))


def ideal_bond_deviation(xyz, neighbours, geometry, bond_radius, metal, ligand, ligands):
    """coordination.py:286-360: |sin| of the angle (and, unless the ligand atom spins freely, the
    dihedral) between the metal-ligand direction and the ligand's free valence."""
    geom = int(geometry[ligand])
    nbs = list(neighbours[ligand])
    length = float(bond_radius[ligand] + bond_radius[metal])
    positions = bond_positions(xyz[ligand], geom, length, [xyz[j] for j in nbs])
    if not positions:
        return 1.0
    rotates = geom == 4 and len(nbs) == 1
    bidentate, mate, shared = False, None, None
    for other in ligands:
        if other != ligand:
            common = sorted(set(nbs) & set(neighbours[other]))
            if common:
                bidentate, mate, shared = True, other, common[0]
                break
    if not nbs:
        return 0.0
    nb = nbs[0]
    nn = next((a for a in neighbours[nb] if a != ligand), None)
    if nn is None:
        rotates = True
    L, Mx, Nb = xyz[ligand], xyz[metal], xyz[nb]
    actual_angle = angle(Nb, L, Mx)
    ideal_angles = [angle(Nb, L, positions[0])]
    ideal_dihedrals = []
    if bidentate and not rotates:
        c = xyz[shared]
        bi = c + 1.5 * ((L - c) + (xyz[mate] - c))
        ideal_angles.append(angle(Nb, L, bi))
        ideal_dihedrals.append(dihedral(xyz[nn], Nb, L, bi))
    out = abs(math.sin(math.radians(min(d - actual_angle for d in ideal_angles))))
    if rotates:
        return out
    actual_dihedral = dihedral(xyz[nn], Nb, L, Mx)
    ideal_dihedrals.append(dihedral(xyz[nn], Nb, L, positions[0]))
    return out + abs(math.sin(math.radians(min(d - actual_dihedral for d in ideal_dihedrals))))


# --------------------------------------------------------------------------------------
# the objective
# --------------------------------------------------------------------------------------
def coordination_sphere(xyz, metal, candidate, radius=3.0, distance=0.0, prevent_intruders=True):
    """coordination.py:233-283: atoms within ``radius`` of the metal (``specifier.zone``:
    d <= radius), walked by increasing |distance - d|; candidate atoms farther than 1.0 on that
    key are kept, and with ``prevent_intruders`` the walk ends at the first atom that is not.
    Returns [(key, atom)]."""
    d = np.sqrt(((xyz - xyz[metal]) ** 2).sum(axis=1))
    zone = [i for i in np.flatnonzero(d <= radius) if i != metal]
    keyed = sorted((abs(distance - float(d[i])), int(i)) for i in zone)
    out = []
    for key, i in keyed:
        if candidate[i] and key > 1.0:
            out.append((key, i))
        elif prevent_intruders:
            break
    return out


def coordination(xyz, metal, candidate, neighbours, geometry, bond_radius, template,
                 radius=3.0, distance=0.0, min_atoms=1, prevent_intruders=True, enforce_all_residues=False,
                 center_of_mass_correction=False, distance_correction=False, weight=-1.0):
    """Coordination.evaluate (coordination.py:180-231).  ``template``: the geometry's vertices with
    the metal's place last.  ``enforce_all_residues`` compares a set with a list
    (coordination.py:264) and is therefore never satisfied: the penalty is returned whenever it is
    set, as the reference does."""
    xyz = np.asarray(xyz, np.float64)
    template = np.asarray(template, np.float64)
    n_vertices = len(template) - 1
    sphere = coordination_sphere(xyz, metal, candidate, radius, distance, prevent_intruders)
    if enforce_all_residues:
        return -1000.0 * weight
    missing = min_atoms - len(sphere)
    if missing > 0:
        return -100.0 * missing * weight
    ligands = [i for _, i in sphere[:n_vertices]]
    points = np.array([xyz[i] for i in ligands] + [xyz[metal]])
    rmsd, ok = rigid_cpd_rmsd(points, template, guess_steps=2, max_iterations=10)
    if not ok:
        return -1000.0 * weight
    score = rmsd + sum(ideal_bond_deviation(xyz, neighbours, geometry, bond_radius, metal, l, ligands) for l in ligands)
    if center_of_mass_correction:
        score += float(np.linalg.norm(points[-1] - points[:-1].mean(axis=0)))
    if distance_correction:
        for l in ligands:
            score += abs(float(np.linalg.norm(xyz[l] - xyz[metal])) - float(bond_radius[l] + bond_radius[metal]))
    return score
