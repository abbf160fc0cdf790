"""Glue between the product's static model (gaudi_b200.model) and the CPU oracle
(oracle/gaudi_oracle.py).  Test infrastructure only.

Besides calling the oracle it computes, per individual, how much of each score hangs on
decisions that lie within ``EPS`` = 1e-4 A of a cutoff (zone radius, DetectClash thresholds,
assumedMaxVdw bound).  north_star exempts exactly those pairs from pair-set identity, so
parity tests allow the CUDA scores to differ from the oracle by at most that slack.
"""
from __future__ import annotations

import os
import sys
from types import SimpleNamespace

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import gaudi_oracle as O  # noqa: E402

EPS = 1e-4
_graph_cache = {}


def torsion_list(mob):
    out, rot = [], []
    for t in range(mob.n_torsions):
        idx = mob.moving_indices(t)
        src = int(mob.tors_src[t])
        if src >= 0:
            out.append((src, (int(mob.tors_fixed[t]), int(mob.tors_moving[t]), idx, bool(mob.tors_snap[t]))))
        else:
            code = -(src + 1)
            rot.append((int(mob.tors_fixed[t]), int(mob.tors_moving[t]), idx, code >> 2, code & 3))
    return out, rot


def express(mob, tors, xform, rot_allele=None):
    """Oracle express of one individual -> (M, 3) fp64 world coordinates."""
    tl, rl = torsion_list(mob)
    torsions = [t for _, t in tl]
    allele = [float(tors[src]) for src, _ in tl] if tors is not None else []
    tables = None
    if rl:
        tables = [mob.rot_delta[mob.rot_ptr[r]:mob.rot_ptr[r + 1]] for r in range(mob.n_rotamer_residues)]
    xmask = (mob.flags & O.F_XFORM) != 0
    return O.express(mob.xyz0, torsions, allele, xform, mob.origin_atom, mob.torsion_before_search,
                     xmask, rl or None, rot_allele, tables)


def _system(mob, rec):
    # keyed by identity; the cached entry holds references to both objects so that their ids
    # cannot be recycled for a different model while the entry is alive
    key = (id(mob), id(rec))
    hit = _graph_cache.get(key)
    if hit is None:
        m = mob.n_atoms
        bonds = [(int(a), int(b)) for a, b in mob.bonds] + [(int(a) + m, int(b) + m) for a, b in rec.bonds]
        extra = getattr(rec, "ms_bonds", None)   # bonds mobile atom - static atom (flexible side chains)
        if extra is not None:
            bonds += [(int(a), int(b) + m) for a, b in extra]
        graph = O.Graph(m + rec.n_atoms, bonds)
        radius = np.concatenate([np.asarray(mob.radius, np.float32), np.asarray(rec.radius, np.float32)]).astype(np.float64)
        flags = np.concatenate([mob.flags, rec.flags]).astype(np.uint8)
        hit = _graph_cache[key] = (graph, radius, flags, mob, rec)
    return hit[:3]


def contacts(mob, rec, params, xyz_mobile, with_slack=True):
    graph, radius, flags = _system(mob, rec)
    coords = np.concatenate([np.asarray(xyz_mobile, np.float64), rec.xyz])
    clash, hydro, neg, pos = O.contacts(
        coords, radius, flags, graph, params.radius, params.clash_threshold,
        params.hydrophobic_threshold, 0.0, params.weight, params.bond_separation,
        params.only_internal, params.hbond_allowance, params.assumed_max_vdw, return_pairs=True,
        residue=(np.concatenate([mob.residue, rec.residue]) if not getattr(params, "same_residue", True) else None))
    res = SimpleNamespace(clash_raw=clash, hydro=hydro, neg=neg, pos=pos, n_clash=len(neg), n_hydro=len(pos),
                          clash_slack=0.0, hydro_slack=0.0, borderline=False)
    res.clash = clash
    if params.cutoff and clash > params.cutoff:
        res.clash = -1000 * params.weight
    if with_slack:
        cs, hs = _slack(coords, radius, flags, graph, params)
        res.clash_slack, res.hydro_slack = cs, hs
        res.borderline = (cs > 0) or (hs > 0)
        if params.cutoff and (abs(clash - params.cutoff) <= cs or res.clash != clash and cs > 0):
            res.clash_slack = float("inf") if abs(clash - params.cutoff) <= cs else cs
    return res


def _slack(coords, radius, flags, graph, p):
    """Sum of the largest possible |contribution| of every pair whose acceptance depends on a
    comparison that is within EPS of equality."""
    from scipy.spatial import cKDTree
    seeds = np.flatnonzero(flags & O.F_SEED)
    n = len(coords)
    if p.only_internal:
        inz = np.zeros(n, bool); inz[seeds] = True
        border = np.zeros(n, bool)
    else:
        st = cKDTree(coords[seeds])
        dmin, _ = st.query(coords, k=1)
        is_seed = np.zeros(n, bool); is_seed[seeds] = True
        inz = is_seed | (dmin <= p.radius + EPS)
        border = ~is_seed & (np.abs(dmin - p.radius) <= EPS)
    atoms = np.flatnonzero(inz)
    if len(atoms) < 2:
        return 0.0, 0.0
    tree = cKDTree(coords[atoms])
    thr_min = min(p.hydrophobic_threshold, p.clash_threshold)
    rmax = float(radius[atoms].max())
    pairs = tree.query_pairs(2 * rmax - thr_min + 2 * EPS, output_type="ndarray")
    cs = hs = 0.0
    for ia, ib in pairs:
        a, b = int(atoms[ia]), int(atoms[ib])
        ra, rb = radius[a], radius[b]
        d = float(np.linalg.norm(coords[a] - coords[b]))
        fa, fb = int(flags[a]), int(flags[b])
        hb = ((fa & O.F_DONOR) and (fb & O.F_ACCEPTOR)) or ((fb & O.F_DONOR) and (fa & O.F_ACCEPTOR))
        ov = ra + rb - d - (p.hbond_allowance if hb else 0.0)
        if ov < thr_min - EPS:
            continue
        near = border[a] or border[b]
        near = near or abs(ov - p.hydrophobic_threshold) <= EPS or abs(ov - p.clash_threshold) <= EPS
        for thr in (p.hydrophobic_threshold, p.clash_threshold):
            near = near or abs(d - (max(ra, rb) + p.assumed_max_vdw - thr)) <= EPS
        if not near:
            continue
        if b in graph.within(a, p.bond_separation):
            continue
        if ov >= p.clash_threshold - EPS:
            cs += 2 * abs(O.vdw_vol_overlap(ra, rb, max(ov, 1e-9)))
        if (fa & O.F_HYDROPHOBIC) and (fb & O.F_HYDROPHOBIC) and ov <= p.clash_threshold + EPS:
            hs += 2 * abs(O.lennard_jones(ra, rb, ov))
    return cs, hs


def rmsd_matrix(xyz, rows=None):
    return O.rmsd_matrix(xyz, rows)


def oracle():
    return O


# --------------------------------------------------------------------------------------
# Coordination objective (oracle/coordination_oracle.py)
# --------------------------------------------------------------------------------------
import coordination_oracle as CO  # noqa: E402


def coordination(site, xyz_world):
    """Oracle score of one individual for a site table (the dict ``Engine.set_coordination`` takes)
    and the world coordinates of every atom of the job."""
    return CO.coordination(xyz_world, int(site["metal"]), np.asarray(site["candidate"], bool), site["neighbours"],
                           site["geometry"], site["bond_radius"], site["template"], radius=site["radius"],
                           distance=site["distance"], min_atoms=site["min_atoms"],
                           prevent_intruders=site.get("prevent_intruders", True),
                           enforce_all_residues=site.get("enforce_all_residues", False),
                           center_of_mass_correction=site.get("center_of_mass_correction", False),
                           distance_correction=site.get("distance_correction", False), weight=site["weight"])


def coordination_world(site, pose_world):
    """World coordinates of all atoms: static ones from the site table, mobile ones from the
    expressed pose (rows ``site['mobile_row']``)."""
    xyz = np.array(site["xyz"], np.float64)
    row = np.asarray(site["mobile_row"])
    xyz[row >= 0] = pose_world[row[row >= 0]]
    return xyz
