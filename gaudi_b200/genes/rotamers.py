"""Rotamers gene (gaudi/genes/rotamers.py): allele = one float in [0, 1) per residue selecting a
row of that residue's chi table; expression turns the chi bonds by ``chi - chi0`` with the CA
side fixed (rotamers.py:141-152,186-210).

The reference pulls the chi tables from Chimera's bundled Dunbrack / Dynameomics libraries and
sorts them by RMSD to the original side chain (rotamers.py:165-184).  Those libraries are not
available here (and their contents are Chimera-version dependent, tests/test_genes_rotamers.py:
42-55), so the table is an *input*: ``library`` is a mapping ``{residue_type: [[chi1, ...], ...]}``
(degrees, already in the order the caller wants); ``with_original`` prepends the residue's own
chis exactly as the reference does (rotamers.py:180-181)."""
from __future__ import annotations

import random

from .. import deaplite, molio
from . import GeneProvider


def enable(**kwargs):
    kwargs = Rotamers.validate(kwargs)
    return Rotamers(**kwargs)


class Rotamers(GeneProvider):
    _validate = {"residues": "required", "library": None, "with_original": None}
    _residues_without_rotamers = set(("ALA", "GLY"))

    def __init__(self, residues=None, library=None, with_original=True, **kwargs):
        GeneProvider.__init__(self, **kwargs)
        self._residues = [tuple(r.split("/")) if isinstance(r, str) else tuple(r) for r in residues]
        self.library = library or {}
        self.with_original = with_original
        self.allele = []
        self.residues = []        # [(molname, position, atom indices, chi quads, chi0, table)]

    def __ready__(self):
        for molname, pos in self._residues:
            gene = self.parent.find_molecule(molname)
            for position in gene.find_residues(pos if pos == "*" else int(pos)):
                atoms = gene.mol.residue_atoms(position)
                resn = gene.mol.resname[atoms[0]][:3].upper()
                quads, chi0 = molio.residue_chis(gene.mol, atoms)
                table = [list(row) for row in self.library.get(resn, [])]
                if self.with_original and chi0:
                    table = [list(chi0)] + table
                if resn in self._residues_without_rotamers or not quads or not table:
                    table = []
                self.residues.append((molname, position, atoms, quads, chi0, table))
                self.allele.append(random.random())

    def mate(self, mate):
        self.allele, mate.allele = deaplite.cx_two_point(self.allele, mate.allele)

    def mutate(self, indpb):
        self.allele = [random.random() if random.random() < indpb else i for i in self.allele]
