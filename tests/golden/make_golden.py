#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ from the reference's own test data.

Run in the build container (where /root/reference exists):

    python tests/golden/make_golden.py

It (1) converts the structure files the reference's hot-path tests use
(/root/reference/tests/data/*) into compact .npz atom tables that travel to the GPU box,
(2) records the golden values those tests assert and (3) imports the
reference's own ``gaudi/similarity.py`` (the only reference module that runs under
Python 3 without UCSF Chimera) to generate RMSD vectors on seeded random inputs.
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


def main():
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


if __name__ == "__main__":
    main()
