"""Load the golden fixtures (tests/golden/*.npz) back into gaudi_b200.molio.Mol objects."""
import json
import os

import numpy as np

from gaudi_b200 import molio

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    mol = molio.Mol([str(x) for x in z["name"]], [str(x) for x in z["resname"]], z["resseq"],
                    [str(x) for x in z["chain"]], [str(x) for x in z["element"]], z["xyz"], z["serial"],
                    hetero=z["hetero"])
    if "sybyl" in z.files:
        mol.sybyl = [str(x) for x in z["sybyl"]]
        mol.bonds = z["bonds"].astype(np.int32)
    else:
        mol.conect = z["conect"].astype(np.int32)
    return mol


def values():
    with open(os.path.join(GOLDEN, "golden_values.json")) as fh:
        return json.load(fh)


def ligand_3pk2():
    """3pk2_ligand.pdb with CONECT bonds and the IDATM proxies of 3pk2_ligand.mol2."""
    lp = load("3pk2_ligand_pdb")
    lp.bonds = lp.conect
    lm = load("3pk2_ligand_mol2")
    ty = dict(zip(lm.name, lm.sybyl))
    lp.sybyl = [ty.get(n, lp.element[i].title()) for i, n in enumerate(lp.name)]
    return lp


def write_mol2(mol, path):
    """A minimal Tripos mol2 of a fixture molecule (what gaudi_b200.molio.read_mol2 reads back):
    lets the tests hand the plugin layer real files, as a job file would."""
    with open(path, "w") as fh:
        fh.write("@<TRIPOS>MOLECULE\nfixture\n%d %d 1\nSMALL\nNO_CHARGES\n\n@<TRIPOS>ATOM\n" % (mol.n_atoms, len(mol.bonds)))
        for i in range(mol.n_atoms):
            fh.write("%7d %-8s %12.6f %12.6f %12.6f %-6s %5d %s%d 0.0000\n" % (
                int(mol.serial[i]), mol.name[i], mol.xyz[i, 0], mol.xyz[i, 1], mol.xyz[i, 2], mol.sybyl[i],
                int(mol.resseq[i]), mol.resname[i], int(mol.resseq[i])))
        fh.write("@<TRIPOS>BOND\n")
        for k, (a, b) in enumerate(mol.bonds):
            fh.write("%6d %6d %6d 1\n" % (k + 1, int(mol.serial[a]), int(mol.serial[b])))
