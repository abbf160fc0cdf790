"""Molecule gene (gaudi/genes/molecule.py): the *source of static topology, coordinates and
radii* for the hot path, and the writer of a result's structure file.  Fragment-library building,
AddH and pdbfixer are out of scope; a Molecule here is one structure
file (PDB with CONECT/distance-perceived bonds, or mol2)."""
from __future__ import annotations

import os

import numpy as np

from .. import molio
from . import GeneProvider


def enable(**kwargs):
    kwargs = Molecule.validate(kwargs)
    return Molecule(**kwargs)


class Molecule(GeneProvider):
    _validate = {"path": "required", "symmetry": None, "hydrogens": None, "pdbfix": None, "vdw_radii": None,
                 "types_from": None, "mol": None}
    _CATALOG = {}

    def __init__(self, path=None, vdw_radii=None, types_from=None, mol=None, **kwargs):
        GeneProvider.__init__(self, **kwargs)
        self.path = path
        self.vdw_radii_overrides = vdw_radii
        key = (os.path.abspath(path) if path else id(mol), types_from)
        if key not in self._CATALOG:           # gaudi/genes/molecule.py:157-177 caches compounds too
            if mol is None:
                ext = os.path.splitext(path)[1].lower()
                if ext == ".mol2":
                    mol = molio.read_mol2(path)
                elif ext == ".pdb":
                    mol = molio.perceive_bonds(molio.read_pdb(path))
                else:
                    raise ValueError("Molecule: unsupported file type %r" % ext)
            if types_from:                     # borrow IDATM proxies from a mol2 of the same molecule
                donor = molio.read_mol2(types_from)
                ty = dict(zip(donor.name, donor.sybyl))
                mol.sybyl = [ty.get(n, mol.element[i].title()) for i, n in enumerate(mol.name)]
            self._CATALOG[key] = mol
        self.mol = self._CATALOG[key]
        self.allele = self.path
        self._expressed_coordinates = None

    # the bits of the reference API the hot path touches -------------------------------
    @property
    def compound(self):
        return self

    def find_atom(self, serial):
        """gaudi/genes/molecule.py:365-371 — index of the atom with this serial number."""
        hits = np.flatnonzero(self.mol.serial == int(serial))
        if not len(hits):
            raise StopIteration("atom serial %s not found in %s" % (serial, self.name))
        return int(hits[0])

    def find_residues(self, position):
        if position == "*":
            return sorted(set(int(r) for r in self.mol.resseq))
        return [int(position)] if int(position) in set(int(r) for r in self.mol.resseq) else []

    def xyz(self):
        """gaudi/genes/molecule.py:415-416: transformed coordinates, float64 (n, 3)."""
        if self._expressed_coordinates is not None:
            return self._expressed_coordinates
        return self.mol.xyz

    def radii(self):
        return molio.vdw_radii(self.mol, self.vdw_radii_overrides)

    def write(self, path=None, name=None, absolute=None, combined_with=None, filetype="mol2"):
        """gaudi/genes/molecule.py:235-261: the molecule with its EXPRESSED coordinates as
        ``<path>/<name>_<gene name>.mol2`` (or ``.pdb``); returns the file name."""
        import tempfile
        if path and name:
            fullname = os.path.join(path, "{}_{}.{}".format(name, self.name, filetype))
        elif absolute:
            fullname = absolute
        else:
            handle, fullname = tempfile.mkstemp(prefix="gaudi", suffix=".{}".format(filetype))
            os.close(handle)
        mol, xyz = self.mol, self.xyz()
        if combined_with:          # the other Molecule genes' atoms in the same file, renumbered (:252-254)
            mol = molio.combine([(self.mol, xyz)] + [(g.mol, g.xyz()) for g in combined_with])
            xyz = mol.xyz
        if filetype == "mol2":
            molio.write_mol2(mol, fullname, xyz=xyz, title=self.name)
        elif filetype == "pdb":
            molio.write_pdb(mol, fullname, xyz=xyz)
        else:
            raise ValueError("Filetype {} not recognized. Try with mol2 or pdb".format(filetype))
        return fullname

    def mutate(self, indpb):
        pass

    def mate(self, gene):
        pass

    def __deepcopy__(self, memo):
        new = self.__class__.__new__(self.__class__)
        new.__dict__.update(self.__dict__)
        return new

    @classmethod
    def clear_cache(cls):
        GeneProvider.clear_cache()
        cls._CATALOG.clear()
