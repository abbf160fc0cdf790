"""Pins the CPU oracle (oracle/gaudi_oracle.py) and the host prep (gaudi_b200.molio) against
every golden value the reference's own hot-path tests assert, using
fixtures generated from /root/reference/tests/data by tests/golden/make_golden.py, and
against outputs of the reference's own gaudi/similarity.py recorded in the build container.
CPU only."""
import numpy as np
import pytest

import golden_util as G
import oracle_adapter as OA
from gaudi_b200 import model as M
from gaudi_b200 import molio

O = OA.oracle()
GV = G.values()


def torsion_list(mol, rot):
    adj = mol.adjacency()
    return [(f, m, np.array(sorted(M.side_of_bond(adj, f, m))), bool(s)) for f, m, s in rot]


def test_G10_atom_counts():
    assert G.load("5dfr_minimized_pdb").n_atoms == GV["G10_atom_counts"]["5dfr_minimized"]
    assert G.load("1amb_pdb").n_atoms == GV["G10_atom_counts"]["1amb"]
    assert G.load("3pk2_ligand_pdb").n_atoms == GV["G10_atom_counts"]["3pk2_ligand"]


def test_G1_backbone_torsion():
    """tests/test_genes_torsion.py:54-63 — pins rotation sign, anchor side and bond order."""
    mol = molio.perceive_bonds(G.load("1amb_pdb"))
    assert len(mol.bonds) == 444
    rot = molio.rotatable_bonds(mol, atom_types=(), atom_names=("CA",))
    assert len(rot) == 82
    assert all("CA" in (mol.name[f], mol.name[m]) for f, m, _ in rot)
    x = O.express(mol.xyz, torsion_list(mol, rot), [90.0] * len(rot))
    g = GV["G1_backbone_torsion_distance"]
    assert abs(np.linalg.norm(x[0] - x[-1]) - g["value"]) < g["tol"]
    # the opposite handedness would not pass
    x = O.express(mol.xyz, torsion_list(mol, rot), [-90.0] * len(rot))
    assert abs(np.linalg.norm(x[0] - x[-1]) - g["value"]) > 1.0


def test_G2_torsion_topology():
    """tests/test_genes_torsion.py:41-51."""
    lig = G.ligand_3pk2()
    g = GV["G2_torsion"]
    assert len(lig.bonds) == g["bonds"]
    rot = molio.rotatable_bonds(lig)
    assert len(rot) == g["rotatable"]
    x = O.express(lig.xyz, torsion_list(lig, rot), [1.0] * len(rot))
    assert abs(np.linalg.norm(x[0] - x[-1]) - g["distance"]) < g["tol"]


@pytest.mark.parametrize("fixture, key, which", [
    ("5dfr_minimized_pdb", "G3_clashes_5dfr", 0),
    ("1amb_pdb", "G4_hydrophobic_1amb", 1),
])
def test_G3_G4_contacts(fixture, key, which):
    """tests/test_objectives_contacts.py:45-50 (the radius there is random: any value works)."""
    mol = molio.perceive_bonds(G.load(fixture))
    rad = molio.vdw_radii(mol)
    flags = molio.atom_flags(mol, seed=True)
    graph = O.Graph(mol.n_atoms, mol.bonds)
    for radius in (0.0, 2.7, 9.9):
        scores = O.contacts(mol.xyz, rad, flags, graph, radius_zone=radius)
        assert abs(scores[which] - GV[key]["value"]) < GV[key]["tol"]


def test_G5_G6_rmsd():
    """tests/test_similarity.py:32-50."""
    a = G.load("3pk2_ligand_pdb").xyz
    b = G.load("3pk2_ligand_rotated_pdb").xyz
    g = GV["G5_rmsd_3pk2"]
    assert abs(O.rmsd_squared(a, b) - g["value"] ** 2) < g["tol_on_square"]
    assert O.rmsd_similar(a, b, g["value"])
    c = G.load("5dfr_minimized_pdb").xyz
    assert O.rmsd_squared(c, c) == 0.0 and O.rmsd_similar(c, c, 0.0)
    assert not O.rmsd_similar(a, c, 100.0)          # different atom counts (similarity.py:65-66)


def test_similarity_matches_reference_module():
    """Outputs of the reference's own gaudi/similarity.py (recorded by make_golden.py)."""
    z = np.load(G.GOLDEN + "/similarity_reference.npz")
    conf, sq, sim, thr = z["conformers"], z["rmsd_squared"], z["similar"], float(z["threshold"])
    got = OA.rmsd_matrix(conf) ** 2
    assert np.allclose(got, sq, rtol=1e-12, atol=1e-12)
    mine = np.array([[O.rmsd_similar(a, b, thr) for b in conf] for a in conf])
    assert np.array_equal(mine, sim)
    assert sim.any() and not sim.all()


def test_G7_chi_convention():
    """tests/test_genes_rotamers.py:44-45."""
    mol = G.load("4c3w_protein_mol2")
    atoms = mol.residue_atoms(5)
    assert mol.resname[atoms[0]] == "ARG"
    quads, chis = molio.residue_chis(mol, atoms)
    g = GV["G7_arg5_chis"]
    assert len(chis) == 4
    # the golden was recorded to 3 decimals
    assert np.abs(np.array(chis) - np.array(g["value"])).max() < 2 * g["tol"]
    # the oracle's dihedral is the same convention
    assert abs(O.dihedral(*[mol.xyz[i] for i in quads[0]]) - chis[0]) < 1e-9


def test_G8_search_property():
    """tests/test_genes_search.py:40-50 — the target atom lands within `radius` of the centre."""
    from gaudi_b200 import synth
    lig = G.load("5er1_ligand_mol2")
    rng = np.random.default_rng(0)
    target = 0   # 'Molecule/1' -> serial 1
    for radius, rotate in ((3.0, False), (15.0, True)):
        t = synth.random_points_in_sphere(8, lig.xyz[target], radius, rng)
        r = synth.random_rotations(8, rng) if rotate else np.repeat(np.eye(3)[None], 8, 0)
        for k in range(8):
            xf = O.multiply_3x4(np.c_[np.eye(3), t[k]], np.c_[r[k], np.zeros(3)])
            x = O.express(lig.xyz, [], [], xf, origin_atom=target)
            assert np.linalg.norm(x[target] - lig.xyz[target]) <= radius + 1e-9
            d0 = np.linalg.norm(lig.xyz[lig.bonds[:, 0]] - lig.xyz[lig.bonds[:, 1]], axis=1)
            d1 = np.linalg.norm(x[lig.bonds[:, 0]] - x[lig.bonds[:, 1]], axis=1)
            assert np.abs(d0 - d1).max() < 1e-9


def test_greedy_niching_restatement():
    """gaudi/algorithms.py:139-155 on a toy case."""
    sim = {(0, 1), (1, 0), (2, 3), (3, 2)}
    fit = {0: 1.0, 1: 2.0, 2: 5.0, 3: 4.0}
    kept = O.greedy_niching([0, 1, 2, 3], lambda i, j: (i, j) in sim, lambda i, j: fit[i] <= fit[j])
    assert kept == [1, 2]


# ---- Coordination objective ------------------------------------------
def coordination_site_4c3w(mobile=(), **kw):
    """Site table of the reference's Coordination test (tests/test_objectives_coordination.py:33-60)
    from the fixture, typed by the product's host prep (molio)."""
    import json
    import os
    from gaudi_b200.objectives.coordination import GEOMETRIES
    mol = G.load("4c3w_mol2")
    g = json.load(open(os.path.join(G.GOLDEN, "coordination_values.json")))["G11_coordination_4c3w"]
    adj = mol.adjacency()
    nbs = [[] for _ in range(mol.n_atoms)]
    for a, b in mol.bonds:
        nbs[int(a)].append(int(b)); nbs[int(b)].append(int(a))
    idatm = [molio.idatm_type(mol, i, adj) for i in range(mol.n_atoms)]
    row = -np.ones(mol.n_atoms, np.int32)
    row[list(mobile)] = np.arange(len(mobile))
    site = {"mobile_row": row, "xyz": mol.xyz.astype(np.float64),
            "candidate": np.array([e.upper() in g["elements"] for e in mol.element], np.uint8),
            "geometry": np.array([molio.IDATM_GEOMETRY.get(t, 3) for t in idatm], np.uint8),
            "bond_radius": np.array([molio.BOND_RADIUS.get(e.upper(), 0.8) for e in mol.element]),
            "neighbours": nbs, "metal": int(np.flatnonzero(mol.serial == g["metal"])[0]),
            "template": np.array(GEOMETRIES[g["geometry"]]), "radius": 3.0, "distance": 0.0, "weight": -1.0,
            "min_atoms": 1}
    site.update(kw)
    return mol, site, g


def test_G11_coordination_goldens():
    """2.4696 and, with both corrections, 3.2405; the six ligand atoms the reference's test names."""
    mol, site, g = coordination_site_4c3w()
    sphere = OA.CO.coordination_sphere(site["xyz"], site["metal"], site["candidate"].astype(bool))[:6]
    assert sorted(int(mol.serial[i]) for _, i in sphere) == sorted(g["ligands"])
    assert abs(OA.coordination(site, site["xyz"]) - g["score"]) < g["tol"]
    site.update(center_of_mass_correction=True, distance_correction=True)
    assert abs(OA.coordination(site, site["xyz"]) - g["score_corrected"]) < g["tol"]
    # penalties of coordination.py:191-199
    site.update(min_atoms=12)
    assert OA.coordination(site, site["xyz"]) == -100.0 * 2 * -1.0        # ten atoms precede the first intruder
    site.update(min_atoms=1, enforce_all_residues=True)
    assert OA.coordination(site, site["xyz"]) == 1000.0


def test_rigid_cpd_matches_reference_module():
    """oracle rigid CPD == the reference's own gaudi/_cpdrift.py (outputs recorded in the build container)."""
    z = np.load(G.GOLDEN + "/cpd_reference.npz")
    names = [str(n) for n in z["template_names"]]
    assert z["ok"].all()
    worst = 0.0
    for pts, n, t, want in zip(z["points"], z["n_points"], z["template"], z["rmsd"]):
        got, ok = OA.CO.rigid_cpd_rmsd(pts[:n], np.array(OA.CO.GEOMETRIES[names[t]]))
        assert ok
        worst = max(worst, abs(got - want))
    assert worst < 1e-9, worst
