"""
Synthetic workloads of the shapes BASELINE.json names.

Seeds are fixed so the GPU run, the CPU oracle and the tests all see the same poses:
ligand seed 1, receptor seed 2, alleles seed 3, RMSD conformers seed 4, rotamers seed 5.
"""
from __future__ import annotations

import numpy as np

from . import model as M

ELEMENT_RADIUS = {"H": 1.0, "C": 1.7, "N": 1.625, "O": 1.5, "S": 1.782}


def _element_flags(elements, h_polar=None):
    flags = np.zeros(len(elements), dtype=np.uint8)
    for i, e in enumerate(elements):
        if e in ("C", "S"):
            flags[i] |= M.F_HYDROPHOBIC
        if e == "N":
            flags[i] |= M.F_DONOR
        if e in ("O", "S"):
            flags[i] |= M.F_ACCEPTOR
        if e == "H" and h_polar is not None and h_polar[i]:
            flags[i] |= M.F_DONOR
    return flags


def make_ligand(n_atoms=40, n_torsions=10, seed=1, bond=1.5):
    """Heavy-atom ligand grown as a random tree around one 6-ring."""
    rng = np.random.default_rng(seed)
    for _attempt in range(200):
        ring_r = bond  # regular hexagon: circumradius == side
        xyz = [np.array([ring_r * np.cos(k * np.pi / 3), ring_r * np.sin(k * np.pi / 3), 0.0])
               for k in range(6)]
        bonds = [(k, (k + 1) % 6) for k in range(6)]
        degree = [2] * 6
        tries = 0
        while len(xyz) < n_atoms and tries < 20000:
            tries += 1
            # favour chain growth so that enough non-terminal acyclic bonds exist
            cand = [i for i in range(len(xyz)) if degree[i] < 3]
            parent = int(rng.choice(cand[-8:])) if rng.random() < 0.7 else int(rng.choice(cand))
            v = rng.normal(size=3)
            v /= np.linalg.norm(v)
            p = xyz[parent] + bond * v
            d = np.linalg.norm(np.asarray(xyz) - p, axis=1)
            d[parent] = 10.0
            if d.min() < 2.1:
                continue
            xyz.append(p)
            bonds.append((parent, len(xyz) - 1))
            degree[parent] += 1
            degree.append(1)
        if len(xyz) < n_atoms:
            continue
        xyz = np.asarray(xyz)
        adj = M.adjacency(n_atoms, bonds)
        # anchor = atom nearest the geometric centre (torsion.py:330-338)
        centre = xyz.mean(axis=0)
        anchor = int(np.argmin(np.linalg.norm(xyz - centre, axis=1)))
        dist = M.graph_distances(adj, anchor)
        rot = []
        for a, b in sorted(bonds, key=lambda ab: min(ab)):
            if a < 6 and b < 6:
                continue                      # ring bond
            if len(adj[a]) < 2 or len(adj[b]) < 2:
                continue                      # terminal
            fixed, moving = (a, b) if dist[a] <= dist[b] else (b, a)
            rot.append((fixed, moving))
        if len(rot) < n_torsions:
            continue
        rot = rot[:n_torsions]
        elements = rng.choice(["C", "N", "O", "S"], size=n_atoms, p=[0.70, 0.12, 0.15, 0.03])
        elements[:6] = "C"
        radius = np.array([ELEMENT_RADIUS[e] for e in elements])
        flags = _element_flags(elements) | np.uint8(M.F_SEED | M.F_XFORM)
        mob = M.build_mobile(xyz, radius, flags, bonds, rot, origin_atom=anchor,
                             torsion_before_search=True)
        mob.elements = list(elements)
        return mob
    raise RuntimeError("could not grow a ligand with %d torsions" % n_torsions)


def make_receptor(n_atoms=5000, seed=2, density=0.085, cavity=8.0, params=None):
    """All-atom receptor box: jittered lattice at protein density with a spherical cavity
    at the origin; "bonds" are pairs closer than 1.9 A."""
    from scipy.spatial import cKDTree
    rng = np.random.default_rng(seed)
    params = params or M.ContactParams()
    a = (1.0 / density) ** (1.0 / 3.0)
    edge = (n_atoms / density + 4.0 / 3.0 * np.pi * cavity ** 3) ** (1.0 / 3.0)
    n = int(np.ceil(edge / a)) + 4
    g = (np.arange(n) - (n - 1) / 2.0) * a
    pts = np.stack(np.meshgrid(g, g, g, indexing="ij"), axis=-1).reshape(-1, 3)
    pts = pts + rng.uniform(-0.5, 0.5, size=pts.shape)
    pts = pts[np.linalg.norm(pts, axis=1) > cavity]
    cheb = np.abs(pts).max(axis=1)
    order = np.argsort(cheb, kind="stable")[:n_atoms]
    pts = pts[np.sort(order)]
    elements = rng.choice(["H", "C", "N", "O", "S"], size=len(pts),
                          p=[0.50, 0.32, 0.08, 0.09, 0.01])
    radius = np.array([ELEMENT_RADIUS[e] for e in elements])
    tree = cKDTree(pts)
    bonds = tree.query_pairs(1.9, output_type="ndarray").astype(np.int32)
    # H is a donor when "bonded" to N/O
    h_polar = np.zeros(len(pts), dtype=bool)
    for x, y in bonds:
        if elements[x] == "H" and elements[y] in ("N", "O"):
            h_polar[x] = True
        if elements[y] == "H" and elements[x] in ("N", "O"):
            h_polar[y] = True
    flags = _element_flags(elements, h_polar)
    rec = M.build_receptor(pts, radius, flags, bonds, params, centre=np.zeros(3))
    rec.elements = list(elements)
    return rec


def random_rotations(n, rng):
    """Uniform random rotation matrices (n, 3, 3) from uniform unit quaternions
    (the role of ``FitMap.search.random_rotation`` at search.py:254)."""
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, x, y, z = q.T
    r = np.empty((n, 3, 3))
    r[:, 0, 0] = 1 - 2 * (y * y + z * z)
    r[:, 0, 1] = 2 * (x * y - z * w)
    r[:, 0, 2] = 2 * (x * z + y * w)
    r[:, 1, 0] = 2 * (x * y + z * w)
    r[:, 1, 1] = 1 - 2 * (x * x + z * z)
    r[:, 1, 2] = 2 * (y * z - x * w)
    r[:, 2, 0] = 2 * (x * z - y * w)
    r[:, 2, 1] = 2 * (y * z + x * w)
    r[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return r


def random_points_in_sphere(n, centre, radius, rng):
    """search.py:311-335 (rejection from the cube), vectorised."""
    out = np.empty((0, 3))
    while len(out) < n:
        p = rng.uniform(-radius, radius, size=(2 * n + 16, 3))
        out = np.concatenate([out, p[(p * p).sum(axis=1) <= radius * radius]])
    return out[:n] + np.asarray(centre, dtype=np.float64)


def make_alleles(pop, n_torsions, seed=3, centre=(0.0, 0.0, 0.0), radius=5.0):
    """Torsion alleles U(-180, 180) rounded to 0.1 deg (torsion.py:205-213) and Search
    alleles as the 3x4 product T.R (search.py:249-255), flattened to (P, 12)."""
    rng = np.random.default_rng(seed)
    tors = np.round(rng.uniform(-180.0, 180.0, size=(pop, n_torsions)), 1)
    rot = random_rotations(pop, rng)
    tr = random_points_in_sphere(pop, centre, radius, rng)
    xf = np.concatenate([rot, tr[:, :, None]], axis=2).reshape(pop, 12)
    return np.ascontiguousarray(tors), np.ascontiguousarray(xf)


def make_conformers(n=16384, n_atoms=60, seed=4):
    """Expressed poses of a 60-atom synthetic ligand: (N, A, 3) float64, built on the host
    with numpy only (used as *input* to the RMSD stage)."""
    lig = make_ligand(n_atoms=n_atoms, n_torsions=10, seed=seed)
    rng = np.random.default_rng(seed)
    rot = random_rotations(n, rng)
    tr = random_points_in_sphere(n, (0, 0, 0), 5.0, rng)
    x0 = lig.xyz0 - lig.xyz0[lig.origin_atom]
    return np.einsum("nij,aj->nai", rot, x0) + tr[:, None, :]
