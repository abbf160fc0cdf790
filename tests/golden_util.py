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


def configs4_job(n_res=12, seed=5, search_radius=3.0):
    """BASELINE.json configs[4] as a job description: Rotamers gene on the ``n_res`` residues with
    side chains nearest to the ligand of the reference's 5er1 test complex (committed fixtures) +
    Torsion + Search, Clashes + Contacts objectives.  The chi tables are seeded random rows (the
    reference takes them from Chimera's Dunbrack library, whose contents are not available here;
    they are inputs to oracle and kernel alike).  Returns ``(cfg, residues)``."""
    from gaudi_b200 import parse
    rng = np.random.default_rng(seed)
    prot, lig = load("5er1_protein_mol2"), load("5er1_ligand_mol2")
    d = np.linalg.norm(prot.xyz[:, None, :] - lig.xyz[None, ::4, :], axis=2).min(axis=1)
    near = []
    for a in np.argsort(d):
        r, nm = int(prot.resseq[a]), prot.resname[a][:3].upper()
        if nm not in ("ALA", "GLY", "PRO") and r not in near:
            near.append(r)
        if len(near) == n_res:
            break
    library = {nm: rng.uniform(-180, 180, size=(int(rng.integers(3, 82)), 4)).tolist()
               for nm in sorted(set(prot.resname[a][:3].upper() for a in range(prot.n_atoms)))}
    cfg = parse.Settings({
        "output": {"verbose": False, "check_every": 0},
        "genes": [{"name": "Protein", "module": "gaudi_b200.genes.molecule", "path": "5er1_protein.mol2", "mol": prot},
                  {"name": "Ligand", "module": "gaudi_b200.genes.molecule", "path": "5er1_ligand.mol2", "mol": lig},
                  {"name": "Rotamers", "module": "gaudi_b200.genes.rotamers",
                   "residues": [("Protein", str(r)) for r in near], "library": library},
                  {"name": "Torsion", "module": "gaudi_b200.genes.torsion", "target": "Ligand"},
                  {"name": "Search", "module": "gaudi_b200.genes.search", "target": "Ligand/1", "radius": search_radius}],
        "objectives": [{"name": "Clashes", "module": "gaudi_b200.objectives.contacts", "which": "clashes",
                        "probes": ["Ligand"], "radius": 5.0, "weight": -1.0},
                       {"name": "Contacts", "module": "gaudi_b200.objectives.contacts", "which": "hydrophobic",
                        "probes": ["Ligand"], "radius": 5.0, "weight": -1.0}]})
    return cfg, near


def configs4_alleles(batch, template, pop, seed=5, search_radius=3.0):
    """Seeded allele arrays for a configs[4] population: torsions (pop, T) rounded to 0.1 degree,
    Search products T.R (pop, 12) with translations in the search sphere, rotamer alleles
    (pop, n_res) in [0, 1)."""
    from gaudi_b200 import synth
    rng = np.random.default_rng(seed + 1000)
    tors = np.round(rng.uniform(-180, 180, size=(pop, batch.n_torsions)), 1)
    centre = np.asarray(template.genes["Search"].center)
    xf = np.concatenate([synth.random_rotations(pop, rng),
                         synth.random_points_in_sphere(pop, centre, search_radius, rng)[:, :, None]],
                        axis=2).reshape(pop, 12)
    rot = rng.random((pop, batch.n_rot_res))
    return tors, xf, rot
