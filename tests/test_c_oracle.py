"""The C restatement of the oracle (oracle/gaudi_oracle.c) against the numpy one
(oracle/gaudi_oracle.py, pinned to the reference's goldens) and against the goldens directly.
CPU only."""
import sys
import os

import numpy as np
import pytest

import golden_util as G
import oracle_adapter as OA
from gaudi_b200 import model as M
from gaudi_b200 import molio, synth

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import c_oracle  # noqa: E402

GV = G.values()


def test_c_oracle_matches_numpy_oracle_synthetic():
    params = M.ContactParams()
    lig = synth.make_ligand(40, 10, seed=1)
    rec = synth.make_receptor(3000, seed=2, params=params)
    tors, xf = synth.make_alleles(300, 10, seed=3)
    co = c_oracle.COracle(lig, rec, params)
    s, c, x = co.evaluate(tors, xf, want_xyz=True)
    for k in range(0, 300, 23):
        ref = OA.express(lig, tors[k], xf[k])
        assert np.abs(ref - x[k]).max() < 1e-10
        r = OA.contacts(lig, rec, params, ref, with_slack=False)
        assert abs(r.clash - s[k, 0]) <= 1e-9 * max(1.0, abs(r.clash))
        assert abs(r.hydro - s[k, 1]) <= 1e-9 * max(1.0, abs(r.hydro))
        assert (r.n_clash, r.n_hydro) == tuple(c[k])
    # single-pose entry point and a small zone radius / penalty
    p2 = M.ContactParams(radius=2.0, cutoff=50.0, weight=-1.0)
    rec2 = M.build_receptor(rec.xyz, rec.radius, rec.flags, rec.bonds, p2, centre=rec.centre)
    co2 = c_oracle.COracle(lig, rec2, p2)
    for k in (0, 5, 9):
        ref = OA.express(lig, tors[k], xf[k])
        r = OA.contacts(lig, rec2, p2, ref, with_slack=False)
        s1, c1 = co2.contacts(ref)
        assert abs(r.clash - s1[0]) <= 1e-9 * max(1.0, abs(r.clash)) and abs(r.hydro - s1[1]) <= 1e-9 * max(1.0, abs(r.hydro))


def test_c_oracle_goldens():
    """Contacts goldens of the reference (tests/test_objectives_contacts.py:45-50) and the torsion
    golden (tests/test_genes_torsion.py:54-63) through the C code."""
    for fixture, key, col in (("5dfr_minimized_pdb", "G3_clashes_5dfr", 0), ("1amb_pdb", "G4_hydrophobic_1amb", 1)):
        mol = molio.perceive_bonds(G.load(fixture))
        for radius in (0.0, 6.1):
            params = M.ContactParams(radius=radius)
            mob = M.build_mobile(mol.xyz, molio.vdw_radii(mol), molio.atom_flags(mol, seed=True, xform=True),
                                 mol.bonds, [])
            co = c_oracle.COracle(mob, M.empty_receptor(), params)
            s, _ = co.contacts(mol.xyz)
            assert abs(s[col] - GV[key]["value"]) < GV[key]["tol"]
    # split the same protein into a mobile and a static part: same score (exclusions across the cut)
    mol = molio.perceive_bonds(G.load("5dfr_minimized_pdb"))
    params = M.ContactParams(radius=4.2)
    mob, rec, _ = M.split_molecule(mol.xyz, molio.vdw_radii(mol), molio.atom_flags(mol, seed=True), mol.bonds,
                                   np.arange(1387, 1460), params)
    s, c = c_oracle.COracle(mob, rec, params).contacts(mob.xyz0)
    assert abs(s[0] - GV["G3_clashes_5dfr"]["value"]) < 1e-4 and c[0] == 5
    mol = molio.perceive_bonds(G.load("1amb_pdb"))
    rot = molio.rotatable_bonds(mol, atom_types=(), atom_names=("CA",))
    mob = M.build_mobile(mol.xyz, molio.vdw_radii(mol), molio.atom_flags(mol, seed=True), mol.bonds,
                         [(f, m) for f, m, _ in rot])
    co = c_oracle.COracle(mob, M.empty_receptor(), M.ContactParams(only_internal=True))
    _, _, x = co.evaluate(np.full((1, len(rot)), 90.0), None, want_xyz=True)
    g = GV["G1_backbone_torsion_distance"]
    assert abs(np.linalg.norm(x[0, 0] - x[0, -1]) - g["value"]) < g["tol"]


def test_c_oracle_rmsd_matches_reference_module_outputs():
    z = np.load(G.GOLDEN + "/similarity_reference.npz")
    lig = synth.make_ligand(12, 2, seed=1)
    co = c_oracle.COracle(lig, M.empty_receptor(), M.ContactParams(only_internal=True))
    got = co.rmsd_matrix(z["conformers"])
    assert np.allclose(got ** 2, z["rmsd_squared"], rtol=1e-12, atol=1e-12)
    assert np.allclose(co.rmsd_matrix(z["conformers"], 5, 17), got[5:17])
