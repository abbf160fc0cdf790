"""Torsion gene (gaudi/genes/torsion.py): allele = list of bond-rotation angles in degrees,
deltas from the input geometry.  Variation (SBX crossover, polynomial mutation, random reset,
rounding to ``precision``) stays on the host exactly as in the reference
(torsion.py:173-213); expression is batched on the GPU (``gmm_express``)."""
from __future__ import annotations

import random

from .. import deaplite, molio
from . import GeneProvider


def enable(**kwargs):
    kwargs = Torsion.validate(kwargs)
    return Torsion(**kwargs)


class Torsion(GeneProvider):
    _validate = {"target": "required", "flexibility": None, "max_bonds": None, "anchor": None,
                 "rotatable_atom_types": None, "rotatable_atom_names": None, "rotatable_elements": None,
                 "non_rotatable_bonds": None, "rotatable_bonds": None, "precision": None}
    BONDS_ROTS = {}

    def __init__(self, target=None, flexibility=360.0, max_bonds=None, anchor=None,
                 rotatable_atom_types=("C3", "N3", "C2", "N2", "P"), rotatable_atom_names=(),
                 rotatable_elements=(), non_rotatable_bonds=(), rotatable_bonds=(), precision=1, **kwargs):
        GeneProvider.__init__(self, **kwargs)
        self.target = target
        self.flexibility = 360.0 if flexibility > 360 else float(flexibility)
        self.max_bonds = max_bonds
        self.rotatable_atom_types = tuple(rotatable_atom_types)
        self.rotatable_atom_names = tuple(rotatable_atom_names)
        self.rotatable_elements = tuple(rotatable_elements)
        self.non_rotatable_bonds = non_rotatable_bonds
        self.concrete_rotatable_bonds = rotatable_bonds
        self.precision = int(precision)
        self._anchor = anchor
        self.allele = [self.random_angle() for _ in range(50)]      # torsion.py:146

    def __expression_hooks__(self):
        if self.max_bonds is None:
            self.max_bonds = len(self.rotatable_bonds)
        self.allele = [self.random_angle() for _ in range(self.max_bonds)]

    @property
    def molecule(self):
        return self.parent.find_molecule(self.target)

    @property
    def anchor(self):
        """torsion.py:309-342: explicit anchor, else the atom nearest the geometric centre."""
        gene = self.molecule
        if self._anchor is not None:
            mol_name, serial = self._anchor
            return self.parent.find_molecule(mol_name).find_atom(serial)
        return molio.nearest_atom_to_centre(gene.mol)

    @property
    def rotatable_bonds(self):
        """List of ``(fixed, moving, snap)`` atom-index triples, cached per molecule like the
        reference caches BondRots on the chimera molecule (torsion.py:239-243)."""
        gene = self.molecule
        key = (id(gene.mol), self.rotatable_atom_types, self.rotatable_atom_names, self.rotatable_elements,
               tuple(map(tuple, self.concrete_rotatable_bonds)), self._anchor and tuple(self._anchor))
        hit = self.BONDS_ROTS.get(key)
        if hit is None:
            explicit = None
            if self.concrete_rotatable_bonds:
                explicit = [(gene.find_atom(a), gene.find_atom(b), gene.find_atom(an))
                            for a, b, an in self.concrete_rotatable_bonds]
            skip = []
            for (ma, sa), (mb, sb) in self.non_rotatable_bonds:
                skip.append((self.parent.find_molecule(ma).find_atom(sa), self.parent.find_molecule(mb).find_atom(sb)))
            hit = self.BONDS_ROTS[key] = molio.rotatable_bonds(
                gene.mol, self.rotatable_atom_types, self.rotatable_atom_names, self.rotatable_elements,
                anchor=self.anchor, explicit=explicit, non_rotatable=skip)
        return hit

    def mate(self, mate):
        a, b = deaplite.cx_simulated_binary_bounded(self.allele, mate.allele, eta=self.cxeta,
                                                    low=-0.5 * self.flexibility, up=0.5 * self.flexibility)
        self.allele[:] = [round(n, self.precision) for n in a]
        mate.allele[:] = [round(n, self.precision) for n in b]

    def mutate(self, indpb):
        if random.random() < 0.5:
            allele, = deaplite.mut_polynomial_bounded(self.allele, indpb=self.indpb, eta=self.mteta,
                                                      low=-0.5 * self.flexibility, up=0.5 * self.flexibility)
            self.allele[:] = [round(n, self.precision) for n in allele]
        else:
            self.allele = [self.random_angle() for _ in range(self.max_bonds)]

    def random_angle(self):
        return round(random.uniform(-0.5 * self.flexibility, 0.5 * self.flexibility), self.precision)

    def clear_cache(self):
        GeneProvider.clear_cache()
        self.BONDS_ROTS.clear()
