"""Minimal molecule reader and typing layer (host prep, no Chimera).

The reference gets all of this from UCSF Chimera when ``gaudi.genes.molecule.Molecule``
opens a file (gaudi/genes/molecule.py:296-362): atoms, bonds, IDATM types, default VdW
radii, H-bond donor/acceptor classification.  Chimera is not available here, so the
pieces the hot path needs are restated from Chimera's documented behaviour; the parts
that the reference's own tests pin are checked in
tests/test_oracle_golden.py.

Nothing here runs per evaluation: it produces the static arrays of ``gaudi_b200.model``.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import model as M

# CSD covalent bond radii used by Chimera's distance-based connectivity (bond when
# d <= r_a + r_b + 0.4 A).
BOND_RADIUS = {
    "H": 0.23, "C": 0.68, "N": 0.68, "O": 0.68, "S": 1.02, "P": 1.05, "F": 0.64, "CL": 0.99,
    "BR": 1.21, "I": 1.40, "SE": 1.22, "B": 0.83, "SI": 1.20,
    "FE": 1.34, "ZN": 1.45, "MG": 1.10, "CA": 0.99, "MN": 1.35, "CU": 1.52, "NI": 1.50,
    "CO": 1.33, "IR": 1.32, "V": 1.33, "NA": 0.97, "K": 1.33, "RU": 1.40, "RH": 1.45,
    "PT": 1.50, "PD": 1.50, "OS": 1.37, "MO": 1.47, "W": 1.37,
}
BOND_TOLERANCE = 0.4
# chimera.Element.mass: standard atomic weights (Distance with center_of_mass); restated, unpinned
ATOMIC_MASS = {
    "H": 1.008, "C": 12.011, "N": 14.007, "O": 15.999, "S": 32.06, "P": 30.974, "F": 18.998, "CL": 35.45,
    "BR": 79.904, "I": 126.904, "SE": 78.971, "B": 10.81, "SI": 28.085, "FE": 55.845, "ZN": 65.38, "MG": 24.305,
    "CA": 40.078, "MN": 54.938, "CU": 63.546, "NI": 58.693, "CO": 58.933, "IR": 192.217, "V": 50.942,
    "NA": 22.990, "K": 39.098, "RU": 101.07, "RH": 102.906, "PT": 195.084, "PD": 106.42, "OS": 190.23,
    "MO": 95.95, "W": 183.84,
}
METALS = {"FE", "ZN", "MG", "CA", "MN", "CU", "NI", "CO", "IR", "V", "NA", "K", "RU", "RH", "PT",
          "PD", "OS", "MO", "W", "LI", "AL", "CD", "HG", "AU", "AG"}

# Chimera default VdW radii, explicit-hydrogen table (the values the reference's Contacts
# goldens exercise).
RADIUS_EXPLICIT_H = {"H": 1.0, "C": 1.7, "N": 1.625, "O": 1.5, "S": 1.782, "P": 1.871, "F": 1.56,
                     "CL": 1.735, "BR": 1.978, "I": 2.094}
RADIUS_CARBOXYLATE_O = 1.48
# United-atom radii (no hydrogens in the structure) -- restated, not pinned by any golden.
RADIUS_UNITED = {"C": 1.88, "N": 1.64, "O": 1.42, "S": 1.77, "P": 1.871}
RADIUS_FALLBACK = 2.0

SYBYL_TO_IDATM = {
    "C.3": "C3", "C.2": "C2", "C.1": "C1", "C.ar": "Car", "C.cat": "C2",
    "N.3": "N3", "N.2": "N2", "N.1": "N1", "N.ar": "Npl", "N.am": "Npl", "N.pl3": "Npl", "N.4": "N3+",
    "O.3": "O3", "O.2": "O2", "O.co2": "O2-", "S.3": "S3", "S.2": "S2", "S.o": "Sxd", "S.o2": "Son",
    "S.O": "Sxd", "S.O2": "Son", "P.3": "P", "H": "H",
}

# chimera.idatm.typeInfo[type].geometry (4 tetrahedral, 3 planar, 2 linear, 1 single, 0 ion), restated
# from Chimera's published IDATM table; types that are not listed count as planar, as
# gaudi/objectives/coordination.py:307-308 does.  The Coordination goldens exercise O2- and Npl.
IDATM_GEOMETRY = {
    "C3": 4, "C2": 3, "Car": 3, "Cac": 3, "C1": 2, "C1-": 2,
    "N3+": 4, "N3": 4, "N2+": 3, "N2": 3, "Npl": 3, "Ng+": 3, "Ntr": 3, "Nox": 4, "N1+": 2, "N1": 2,
    "O3": 4, "O2": 3, "Oar": 3, "Oar+": 3, "O3-": 4, "O2-": 3, "O1": 2, "O1+": 2,
    "S3+": 4, "S3": 4, "S2": 3, "Sar": 3, "Sac": 4, "Son": 4, "Sxd": 4, "S3-": 4,
    "Pac": 4, "Pox": 4, "P3+": 4, "P": 4, "B": 3, "Bac": 4, "Box": 3,
    "H": 1, "HC": 1, "D": 1, "DC": 1, "F": 1, "Cl": 1, "Br": 1, "I": 1,
}


@dataclass
class Mol:
    name: list            # atom names
    resname: list
    resseq: np.ndarray    # (n,) int
    chain: list
    element: list         # upper-case symbols
    xyz: np.ndarray       # (n, 3) float64
    serial: np.ndarray    # (n,) int
    bonds: np.ndarray = field(default_factory=lambda: np.zeros((0, 2), np.int32))
    sybyl: list = None    # mol2 atom types when read from mol2
    hetero: np.ndarray = None

    @property
    def n_atoms(self):
        return len(self.name)

    def adjacency(self):
        return M.adjacency(self.n_atoms, self.bonds)

    def residue_atoms(self, resseq, chain=None):
        return [i for i in range(self.n_atoms)
                if self.resseq[i] == resseq and (chain is None or self.chain[i] == chain)]


_TWO_LETTER = set(k for k in BOND_RADIUS if len(k) == 2) | METALS


def _element_from_name(field4, hetero):
    """Element from the 4-character PDB atom-name field when columns 77-78 are empty."""
    raw = field4
    s = raw.strip().upper()
    letters = "".join(ch for ch in s if ch.isalpha())
    if not letters:
        return "X"
    if hetero and raw[:2].strip().upper() in _TWO_LETTER and raw[0] != " ":
        return raw[:2].strip().upper()
    # names like 1HB2 / HB2 / HT1 are hydrogens
    if s[0].isdigit() and len(s) > 1 and s[1] == "H":
        return "H"
    return letters[0]


def read_pdb(path):
    names, resn, resq, chain, elem, xyz, serial, het = [], [], [], [], [], [], [], []
    conect = []
    with open(path) as fh:
        for line in fh:
            rec = line[:6]
            if rec in ("ATOM  ", "HETATM"):
                hetero = rec == "HETATM"
                names.append(line[12:16].strip())
                resn.append(line[17:21].strip())
                chain.append(line[21].strip())
                try:
                    resq.append(int(line[22:26]))
                except ValueError:
                    resq.append(0)
                xyz.append((float(line[30:38]), float(line[38:46]), float(line[46:54])))
                e = line[76:78].strip().upper() if len(line) >= 78 else ""
                if not e or not e.isalpha():
                    e = _element_from_name(line[12:16], hetero)
                elem.append(e)
                serial.append(int(line[6:11]))
                het.append(hetero)
            elif rec == "CONECT":
                nums = [line[k:k + 5] for k in range(6, len(line.rstrip("\n")), 5)]
                nums = [int(x) for x in nums if x.strip()]
                if nums:
                    conect.append(nums)
            elif rec == "ENDMDL":
                break
    mol = Mol(names, resn, np.array(resq, np.int64), chain, elem, np.array(xyz, np.float64),
              np.array(serial, np.int64), hetero=np.array(het, bool))
    if conect:
        idx = {int(s): i for i, s in enumerate(mol.serial)}
        pairs = set()
        for nums in conect:
            a = idx.get(nums[0])
            for b in nums[1:]:
                b = idx.get(b)
                if a is not None and b is not None and a != b:
                    pairs.add((min(a, b), max(a, b)))
        mol.conect = np.array(sorted(pairs), np.int32).reshape(-1, 2)
    else:
        mol.conect = np.zeros((0, 2), np.int32)
    return mol


def read_mol2(path):
    names, resn, resq, elem, xyz, serial, sybyl = [], [], [], [], [], [], []
    bonds = []
    section = None
    with open(path) as fh:
        for line in fh:
            s = line.strip()
            if s.startswith("@<TRIPOS>"):
                section = s[9:]
                continue
            if not s or s.startswith("#"):
                continue
            if section == "ATOM":
                f = s.split()
                serial.append(int(f[0]))
                names.append(f[1])
                xyz.append((float(f[2]), float(f[3]), float(f[4])))
                sybyl.append(f[5])
                resq.append(int(f[6]) if len(f) > 6 else 1)
                sub = f[7] if len(f) > 7 else "UNK"
                resn.append("".join(ch for ch in sub if not ch.isdigit()) or sub)
                e = f[5].split(".")[0].upper()
                elem.append(e)
            elif section == "BOND":
                f = s.split()
                bonds.append((int(f[1]), int(f[2])))
    mol = Mol(names, resn, np.array(resq, np.int64), [""] * len(names), elem,
              np.array(xyz, np.float64), np.array(serial, np.int64), sybyl=sybyl,
              hetero=np.zeros(len(names), bool))
    idx = {int(s): i for i, s in enumerate(mol.serial)}
    mol.bonds = np.array(sorted({(min(idx[a], idx[b]), max(idx[a], idx[b])) for a, b in bonds}),
                         np.int32).reshape(-1, 2)
    return mol


def write_mol2(mol, path, xyz=None, title="gaudi_b200"):
    """Tripos mol2 of ``mol`` with coordinates ``xyz`` (default: the molecule's own): what the
    reference gets from Chimera's ``writeMol2`` at gaudi/genes/molecule.py:256 as far as
    ``read_mol2`` and any mol2 reader care — atom ids, names, coordinates, Sybyl types, residue
    numbers and names, bonds.  Partial charges are written as 0 (the hot path never holds any)."""
    xyz = mol.xyz if xyz is None else np.asarray(xyz, np.float64)
    sybyl = mol.sybyl if getattr(mol, "sybyl", None) else [e.title() for e in mol.element]
    bonds = getattr(mol, "bonds", None)
    bonds = np.zeros((0, 2), np.int32) if bonds is None else np.asarray(bonds).reshape(-1, 2)
    with open(path, "w") as fh:
        fh.write("@<TRIPOS>MOLECULE\n%s\n%d %d 1\nSMALL\nNO_CHARGES\n\n@<TRIPOS>ATOM\n" % (title, mol.n_atoms, len(bonds)))
        for i in range(mol.n_atoms):
            fh.write("%7d %-8s %12.6f %12.6f %12.6f %-6s %5d %s%d 0.0000\n" % (
                int(mol.serial[i]), mol.name[i], xyz[i, 0], xyz[i, 1], xyz[i, 2], sybyl[i],
                int(mol.resseq[i]), mol.resname[i], int(mol.resseq[i])))
        fh.write("@<TRIPOS>BOND\n")
        for k, (a, b) in enumerate(bonds):
            fh.write("%6d %6d %6d 1\n" % (k + 1, int(mol.serial[a]), int(mol.serial[b])))
    return path


def combine(parts):
    """One ``Mol`` out of several ``(mol, xyz)`` parts, atoms renumbered 1..N in order, bonds carried
    over: what Chimera's ``writeMol2(..., multimodelHandling='combined')`` / ``pdbWrite`` of a list of
    molecules put into one file (gaudi/genes/molecule.py:252-258)."""
    name, resname, resseq, chain, element, xyz, hetero, sybyl, bonds = [], [], [], [], [], [], [], [], []
    off = 0
    for mol, x in parts:
        x = mol.xyz if x is None else np.asarray(x, np.float64)
        name += list(mol.name); resname += list(mol.resname); resseq += [int(r) for r in mol.resseq]
        chain += list(mol.chain); element += list(mol.element); xyz.append(x)
        hetero += list(mol.hetero) if getattr(mol, "hetero", None) is not None else [False] * mol.n_atoms
        sybyl += list(mol.sybyl) if getattr(mol, "sybyl", None) else [e.title() for e in mol.element]
        b = getattr(mol, "bonds", None)
        if b is not None and len(b):
            bonds.append(np.asarray(b, np.int64).reshape(-1, 2) + off)
        off += mol.n_atoms
    out = Mol(name, resname, np.asarray(resseq), chain, element, np.concatenate(xyz), np.arange(1, off + 1),
              hetero=np.asarray(hetero, bool))
    out.sybyl = sybyl
    out.bonds = (np.concatenate(bonds) if bonds else np.zeros((0, 2), np.int64)).astype(np.int32)
    return out


def write_pdb(mol, path, xyz=None):
    """Minimal PDB (ATOM/HETATM records) of ``mol`` with coordinates ``xyz``."""
    xyz = mol.xyz if xyz is None else np.asarray(xyz, np.float64)
    with open(path, "w") as fh:
        for i in range(mol.n_atoms):
            rec = "HETATM" if (getattr(mol, "hetero", None) is not None and mol.hetero[i]) else "ATOM  "
            fh.write("%s%5d %-4s %3s %1s%4d    %8.3f%8.3f%8.3f  1.00  0.00          %2s\n" % (
                rec, int(mol.serial[i]) % 100000, mol.name[i][:4], mol.resname[i][:3], (mol.chain[i] or " ")[:1],
                int(mol.resseq[i]) % 10000, xyz[i, 0], xyz[i, 1], xyz[i, 2], mol.element[i].title()[:2]))
        fh.write("END\n")
    return path


def perceive_bonds(mol, use_conect=True):
    """Distance-based connectivity (d <= r_a + r_b + 0.4), merged with CONECT records.
    A hydrogen keeps only its shortest bond."""
    from scipy.spatial import cKDTree
    n = mol.n_atoms
    rad = np.array([BOND_RADIUS.get(e, 0.8) for e in mol.element])
    pairs = set()
    if n > 1:
        tree = cKDTree(mol.xyz)
        cand = tree.query_pairs(2 * rad.max() + BOND_TOLERANCE, output_type="ndarray")
        d = np.linalg.norm(mol.xyz[cand[:, 0]] - mol.xyz[cand[:, 1]], axis=1)
        ok = d <= rad[cand[:, 0]] + rad[cand[:, 1]] + BOND_TOLERANCE
        best_h = {}
        for (a, b), dist in zip(cand[ok], d[ok]):
            a, b = int(a), int(b)
            ea, eb = mol.element[a], mol.element[b]
            if ea == "H" and eb == "H":
                continue
            if ea == "H" or eb == "H":
                h = a if ea == "H" else b
                if h not in best_h or dist < best_h[h][0]:
                    best_h[h] = (dist, (min(a, b), max(a, b)))
                continue
            pairs.add((min(a, b), max(a, b)))
        for _, pr in best_h.values():
            pairs.add(pr)
    if use_conect and getattr(mol, "conect", None) is not None:
        for a, b in mol.conect:
            pairs.add((int(a), int(b)))
    mol.bonds = np.array(sorted(pairs), np.int32).reshape(-1, 2)
    return mol


# ---------------------------------------------------------------------------------------
# typing
# ---------------------------------------------------------------------------------------
def carboxylate_oxygens(mol, adj):
    """Oxygens of a deprotonated carboxylate (IDATM O2-): both O on a C are terminal and
    carry no hydrogen."""
    out = set()
    for c in range(mol.n_atoms):
        if mol.element[c] != "C":
            continue
        oxy = [o for o in adj[c] if mol.element[o] == "O" and len(adj[o]) == 1]
        if len(oxy) == 2:
            out.update(oxy)
    return out


def vdw_radii(mol, overrides=None):
    """Chimera default radii.  ``overrides``: element -> radius (``vdw_radii`` key of the
    Molecule gene, gaudi/genes/molecule.py:825-828)."""
    adj = mol.adjacency()
    has_h = any(e == "H" for e in mol.element)
    table = RADIUS_EXPLICIT_H if has_h else dict(RADIUS_EXPLICIT_H, **RADIUS_UNITED)
    carbox = carboxylate_oxygens(mol, adj) if has_h else set()
    rad = np.empty(mol.n_atoms)
    for i, e in enumerate(mol.element):
        r = table.get(e, RADIUS_FALLBACK)
        if i in carbox:
            r = RADIUS_CARBOXYLATE_O
        if overrides and e.title() in overrides:
            r = overrides[e.title()]
        rad[i] = r
    return rad


def hbond_flags(mol):
    """Donor / acceptor bits as DetectClash's hbondAllowance uses them:
    donor = H on N/O/S, or N/O/S carrying an H; acceptor = IDATM substituents < geometry,
    i.e. every O, every S with <= 2 bonds, N with <= 2 bonds (N2/N1) or an sp3 amine N3."""
    adj = mol.adjacency()
    flags = np.zeros(mol.n_atoms, np.uint8)
    for i, e in enumerate(mol.element):
        if e == "H":
            if adj[i] and mol.element[adj[i][0]] in ("N", "O", "S"):
                flags[i] |= M.F_DONOR
        elif e in ("N", "O", "S"):
            if any(mol.element[j] == "H" for j in adj[i]):
                flags[i] |= M.F_DONOR
            if e == "O":
                flags[i] |= M.F_ACCEPTOR
            elif e == "S":
                if len(adj[i]) <= 2:
                    flags[i] |= M.F_ACCEPTOR
            else:
                idatm = idatm_type(mol, i, adj)
                if idatm in ("N2", "N1", "N3"):
                    flags[i] |= M.F_ACCEPTOR
    return flags


def idatm_type(mol, i, adj=None):
    """IDATM-lite: the mol2 type when there is one, else a guess from element and bond count
    (only the distinctions the hot path needs: C3/C2/Car, N3/N3+/N2/Npl, O*, S*, P)."""
    if mol.sybyl is not None:
        return SYBYL_TO_IDATM.get(mol.sybyl[i], mol.sybyl[i].replace(".", ""))
    adj = adj or mol.adjacency()
    e, nb = mol.element[i], len(adj[i])
    if e == "C":
        return "C3" if nb >= 4 else ("C2" if nb == 3 else "C1")
    if e == "N":
        if nb >= 4:
            return "N3+"
        if nb == 3:
            # planar (amide, aromatic, guanidinium) vs pyramidal amine: sum of angles
            v = [mol.xyz[j] - mol.xyz[i] for j in adj[i]]
            v = [x / np.linalg.norm(x) for x in v]
            ang = sum(np.degrees(np.arccos(np.clip(np.dot(v[a], v[b]), -1, 1)))
                      for a, b in ((0, 1), (0, 2), (1, 2)))
            return "Npl" if ang > 345.0 else "N3"
        return "N2" if nb == 2 else "N1"
    if e == "O":
        return "O3" if nb >= 2 else "O2"
    if e == "S":
        return "S3" if nb <= 2 else "Son"
    if e == "P":
        return "P"
    return e.title()


def atom_flags(mol, hydrophobic_elements=("C", "S"), seed=False, xform=False):
    flags = hbond_flags(mol)
    hyd = {e.upper() for e in hydrophobic_elements}
    for i, e in enumerate(mol.element):
        if e in hyd:
            flags[i] |= M.F_HYDROPHOBIC
    if seed:
        flags |= np.uint8(M.F_SEED)
    if xform:
        flags |= np.uint8(M.F_XFORM)
    return flags


# ---------------------------------------------------------------------------------------
# rotatable bonds (gaudi/genes/torsion.py:245-342, gaudi/box.py:246-254)
# ---------------------------------------------------------------------------------------
def nearest_atom_to_centre(mol):
    """torsion.py:345-363 / search.py:362-383: atom nearest the geometric centre."""
    c = mol.xyz.mean(axis=0)
    return int(np.argmin(np.linalg.norm(mol.xyz - c, axis=1)))


def rotatable_bonds(mol, atom_types=("C3", "N3", "C2", "N2", "P"), atom_names=(), elements=(),
                    anchor=None, explicit=None, non_rotatable=()):
    """List of ``(fixed, moving, snap)`` in the reference's order.

    Rule (torsion.py:281-293): both atoms have more than one bond and are not metals, and at
    least one of them has an IDATM type in ``atom_types``, or name ``DUM``, or a listed name
    or element; bonds sorted by the smaller serial number (torsion.py:246-248); bonds in
    cycles are dropped (``chimera.BondRot`` raises, torsion.py:299-304).  The fixed atom is
    the bond atom graph-closer to the anchor (``box.find_nearest``, torsion.py:305).
    ``snap`` marks C2/N2-C2/N2 bonds (torsion.py:159-160).
    """
    adj = mol.adjacency()
    anchor = nearest_atom_to_centre(mol) if anchor is None else int(anchor)
    dist = M.graph_distances(adj, anchor)
    types = [idatm_type(mol, i, adj) for i in range(mol.n_atoms)]
    out = []
    if explicit:
        cand = [(int(a), int(b), int(an)) for a, b, an in explicit]
    else:
        order = sorted(range(len(mol.bonds)), key=lambda k: min(mol.serial[mol.bonds[k][0]],
                                                                mol.serial[mol.bonds[k][1]]))
        cand = [(int(mol.bonds[k][0]), int(mol.bonds[k][1]), None) for k in order]
    skip = {(min(a, b), max(a, b)) for a, b in non_rotatable}
    for a, b, an in cand:
        if (min(a, b), max(a, b)) in skip:
            continue
        if not explicit:
            if len(adj[a]) == 1 or len(adj[b]) == 1 or mol.element[a] in METALS or mol.element[b] in METALS:
                continue
            if not any(mol.name[x] == "DUM" or types[x] in atom_types or mol.name[x] in atom_names
                       or mol.element[x].title() in elements for x in (a, b)):
                continue
        if an is None:
            fixed, moving = (a, b) if dist.get(a, 1 << 30) <= dist.get(b, 1 << 30) else (b, a)
        else:
            d2 = M.graph_distances(adj, an)
            fixed, moving = (a, b) if d2.get(a, 1 << 30) <= d2.get(b, 1 << 30) else (b, a)
        try:
            M.side_of_bond(adj, fixed, moving)
        except ValueError:
            continue        # bond in a cycle
        snap = types[a] in ("C2", "N2") and types[b] in ("C2", "N2")
        out.append((fixed, moving, int(snap)))
    return out


CHI_ATOMS = {
    # standard chi definitions (Chimera phipsi.chiAtoms): residue -> list of 4-atom names
    "ARG": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "CD"), ("CB", "CG", "CD", "NE"), ("CG", "CD", "NE", "CZ")],
    "ASN": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "OD1")],
    "ASP": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "OD1")],
    "CYS": [("N", "CA", "CB", "SG")],
    "GLN": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "CD"), ("CB", "CG", "CD", "OE1")],
    "GLU": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "CD"), ("CB", "CG", "CD", "OE1")],
    "HIS": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "ND1")],
    "ILE": [("N", "CA", "CB", "CG1"), ("CA", "CB", "CG1", "CD1")],
    "LEU": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "CD1")],
    "LYS": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "CD"), ("CB", "CG", "CD", "CE"), ("CG", "CD", "CE", "NZ")],
    "MET": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "SD"), ("CB", "CG", "SD", "CE")],
    "PHE": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "CD1")],
    "PRO": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "CD")],
    "SER": [("N", "CA", "CB", "OG")],
    "THR": [("N", "CA", "CB", "OG1")],
    "TRP": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "CD1")],
    "TYR": [("N", "CA", "CB", "CG"), ("CA", "CB", "CG", "CD1")],
    "VAL": [("N", "CA", "CB", "CG1")],
}


def dihedral(p0, p1, p2, p3):
    """IUPAC dihedral, degrees (the convention of ``chimera.dihedral``)."""
    b0 = p0 - p1
    b1 = p2 - p1
    b2 = p3 - p2
    b1 = b1 / np.linalg.norm(b1)
    v = b0 - np.dot(b0, b1) * b1
    w = b2 - np.dot(b2, b1) * b1
    return float(np.degrees(np.arctan2(np.dot(np.cross(b1, v), w), np.dot(v, w))))


def residue_chis(mol, atoms):
    """(chi atom index quadruples, chi values) of one residue given its atom indices
    (rotamers.py:196-203)."""
    by_name = {mol.name[i]: i for i in atoms}
    resn = mol.resname[atoms[0]][:3].upper()
    quads, values = [], []
    for names in CHI_ATOMS.get(resn, []):
        if not all(nm in by_name for nm in names):
            break
        q = tuple(by_name[nm] for nm in names)
        quads.append(q)
        values.append(dihedral(*[mol.xyz[i] for i in q]))
    return quads, values
