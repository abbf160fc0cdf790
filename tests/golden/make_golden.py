#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ from the reference's own test data.

Run in the build container (where /root/reference exists):

    python tests/golden/make_golden.py

It (1) converts the structure files the reference's hot-path tests use
(/root/reference/tests/data/*) into compact .npz atom tables that travel to the GPU box,
(2) records the golden values those tests assert and (3) imports the
reference's own ``gaudi/similarity.py`` and ``gaudi/_cpdrift.py`` (the reference modules that run
under Python 3 without UCSF Chimera) to generate RMSD vectors and rigid-CPD results on seeded
random inputs.

    python tests/golden/make_golden.py coordination      # only the Coordination fixtures (added later;
                                                         # leaves the other files byte-identical)
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"
DATA = os.path.join(REF, "tests", "data")

from gaudi_b200 import molio  # noqa: E402


def dump(mol, out):
    kw = dict(name=np.array(mol.name), resname=np.array(mol.resname), resseq=mol.resseq,
              chain=np.array(mol.chain), element=np.array(mol.element), xyz=mol.xyz, serial=mol.serial,
              hetero=mol.hetero)
    if mol.sybyl is not None:
        kw["sybyl"] = np.array(mol.sybyl)
        kw["bonds"] = mol.bonds
    else:
        kw["conect"] = mol.conect
    np.savez_compressed(os.path.join(HERE, out), **kw)


def coordination_fixtures():
    """4c3w.mol2 (the complex of tests/test_objectives_coordination.py) and the reference's own
    rigid CPD (gaudi/_cpdrift.py, loaded by path: the package's __init__ imports Chimera) on
    seeded point sets: noisy, rotated, scaled picks of the geometry templates' vertices."""
    import importlib.util
    dump(molio.read_mol2(os.path.join(DATA, "4c3w.mol2")), "4c3w_mol2.npz")
    spec = importlib.util.spec_from_file_location("ref_cpdrift", os.path.join(REF, "gaudi", "_cpdrift.py"))
    cpd = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(cpd)
    sys.path.insert(0, os.path.join(ROOT))
    from oracle.coordination_oracle import GEOMETRIES
    rng = np.random.default_rng(20261107)
    names = sorted(GEOMETRIES)
    X = np.zeros((240, 9, 3)); n_pts = np.zeros(240, np.int32); which = np.zeros(240, np.int32)
    rmsd = np.zeros(240); ok = np.zeros(240, bool)
    for k in range(240):
        which[k] = k % len(names)
        Y = np.array(GEOMETRIES[names[which[k]]])
        n = int(rng.integers(2, len(Y) + 1))
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        w, x, y, z = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                      [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                      [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
        pick = list(rng.permutation(len(Y) - 1)[:n - 1]) + [len(Y) - 1]
        pts = 2.1 * Y[pick].dot(R.T) + rng.normal(size=3) * 5 + rng.normal(scale=rng.choice([0.0, 0.05, 0.3, 1.0]), size=(n, 3))
        X[k, :n] = pts; n_pts[k] = n
        try:
            rmsd[k] = cpd.coherent_point_drift(pts, Y.copy(), method="rigid", guess_steps=2, max_iterations=10)[2]
            ok[k] = True
        except Exception:
            rmsd[k] = np.nan
    np.savez_compressed(os.path.join(HERE, "cpd_reference.npz"), points=X, n_points=n_pts, template=which,
                        template_names=np.array(names), rmsd=rmsd, ok=ok)
    with open(os.path.join(HERE, "coordination_values.json"), "w") as fh:
        json.dump({"G11_coordination_4c3w": {"metal": 1005, "ligands": [1013, 1014, 1021, 1023, 1024, 408],
                                             "geometry": "octahedron", "elements": ["C", "N", "O", "S"],
                                             "score": 2.4696, "score_corrected": 3.2405, "tol": 1e-4,
                                             "source": "tests/test_objectives_coordination.py:33-60"}},
                  fh, indent=1, sort_keys=True)
    print("wrote the Coordination fixtures to", HERE, "(CPD cases that failed in the reference: %d)" % int((~ok).sum()))


def main():
    if sys.argv[1:] == ["coordination"]:
        return coordination_fixtures()
    for f in ("1amb.pdb", "5dfr_minimized.pdb", "3pk2_ligand.pdb", "3pk2_ligand_rotated.pdb"):
        dump(molio.read_pdb(os.path.join(DATA, f)), f.replace(".pdb", "_pdb.npz"))
    for f in ("3pk2_ligand.mol2", "4c3w_protein.mol2", "5er1_ligand.mol2", "5er1_protein.mol2"):
        dump(molio.read_mol2(os.path.join(DATA, f)), f.replace(".mol2", "_mol2.npz"))

    # values asserted by the reference's tests (file:line in the reference repository)
    golden = {
        "G1_backbone_torsion_distance": {"value": 58.986905376603815, "tol": 1e-4,
                                         "source": "tests/test_genes_torsion.py:54-63"},
        "G2_torsion": {"bonds": 49, "rotatable": 8, "distance": 3.796920067633768, "tol": 1e-4,
                       "source": "tests/test_genes_torsion.py:41-51"},
        "G3_clashes_5dfr": {"value": 6.7255, "tol": 1e-4, "source": "tests/test_objectives_contacts.py:45-50"},
        "G4_hydrophobic_1amb": {"value": -48.9072, "tol": 1e-4, "source": "tests/test_objectives_contacts.py:45-50"},
        "G5_rmsd_3pk2": {"value": 5.911, "tol_on_square": 0.01, "source": "tests/test_similarity.py:32-50"},
        "G6_rmsd_same": {"value": 0.0, "tol_on_square": 0.01, "source": "tests/test_similarity.py:33"},
        "G7_arg5_chis": {"value": [179.734, 178.061, 60.608, 90.076], "tol": 1e-3,
                         "source": "tests/test_genes_rotamers.py:44-45,70-71"},
        "G10_atom_counts": {"5dfr_minimized": 2489, "1amb": 438, "3pk2_ligand": 41,
                            "source": "tests/test_genes_molecule.py:31-35"},
    }
    with open(os.path.join(HERE, "golden_values.json"), "w") as fh:
        json.dump(golden, fh, indent=1, sort_keys=True)

    # the reference's own similarity module, run here on seeded inputs
    sys.path.insert(0, REF)
    from gaudi.similarity import _rmsd_squared, rmsd

    class _Mol(object):
        def __init__(self, c):
            self._expressed_coordinates = c

    class _Ind(object):
        def __init__(self, c):
            self.m = _Mol(c)

        def find_molecule(self, name):
            return self.m

    rng = np.random.default_rng(20261106)
    conf = rng.normal(scale=6.0, size=(48, 60, 3))
    conf[1::2] = conf[0::2] + rng.normal(scale=1.4, size=(24, 60, 3))   # pairs around the threshold
    sq = np.array([[_rmsd_squared(a, b) for b in conf] for a in conf])
    sim = np.array([[rmsd(_Ind(a), _Ind(b), ["Ligand"], 2.5) for b in conf] for a in conf])
    np.savez_compressed(os.path.join(HERE, "similarity_reference.npz"), conformers=conf, rmsd_squared=sq,
                        similar=sim, threshold=2.5)
    print("wrote fixtures to", HERE)
    coordination_fixtures()


if __name__ == "__main__":
    main()
