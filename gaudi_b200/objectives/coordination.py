"""Coordination objective (gaudi/objectives/coordination.py): how well the atoms around a metal
centre match a coordination geometry.  Same YAML keys, defaults and penalties as the reference;
the whole population is scored by ``gmm_coordination`` (sphere search, rigid coherent point drift
onto the geometry's vertices, directionality of the ligand atoms' free valences, optional
centre-of-mass and bond-length corrections)."""
from __future__ import annotations

import numpy as np

from . import ObjectiveProvider
from .. import molio
from .distance import _spec

# coordination.py:51-64 (the metal's place is the last point)
GEOMETRIES = {
    "cube": [(1.0, 1.0, 1.0), (1.0, 1.0, -1.0), (1.0, -1.0, 1.0), (1.0, -1.0, -1.0), (-1.0, 1.0, 1.0), (-1.0, 1.0, -1.0),
             (-1.0, -1.0, 1.0), (-1.0, -1.0, -1.0), (0.0, 0.0, 0.0)],
    "hexagonal bipyramid": [(0.0, 1.0, 0.0), (0.0, -1.0, 0.0), (0.0, 0.0, 1.0), (0.0, 0.0, -1.0), (0.8660, 0.0, 0.5),
                            (0.8660, 0.0, -0.5), (-0.8660, 0.0, 0.5), (-0.8660, 0.0, -0.5), (0.0, 0.0, 0.0)],
    "linear": [(0.0, 1.0, 0.0), (0.0, -1.0, 0.0), (0.0, 0.0, 0.0)],
    "octahedron": [(1.0, 0.0, 0.0), (-1.0, 0.0, 0.0), (0.0, 1.0, 0.0), (0.0, -1.0, 0.0), (0.0, 0.0, 1.0),
                   (0.0, 0.0, -1.0), (0.0, 0.0, 0.0)],
    "pentagonal bipyramid": [(0.0, 1.0, 0.0), (0.0, -1.0, 0.0), (0.0, 0.0, 1.0), (0.9511, 0.0, 0.30901),
                             (0.5878, 0.0, -0.8090), (-0.5878, 0.0, -0.8090), (-0.9511, 0.0, 0.3090), (0.0, 0.0, 0.0)],
    "square planar": [(1.0, 0.0, 0.0), (-1.0, 0.0, 0.0), (0.0, 1.0, 0.0), (0.0, -1.0, 0.0), (0.0, 0.0, 0.0)],
    "square pyramid": [(0.0, 1.0, 0.0), (-0.6123, -0.5, 0.6123), (0.6123, -0.5, 0.6123), (-0.6123, -0.5, -0.6123),
                       (0.6123, -0.5, -0.6123), (0.0, 0.0, 0.0)],
    "tetrahedral": [(1.0, 1.0, 1.0), (-1.0, -1.0, 1.0), (-1.0, 1.0, -1.0), (1.0, -1.0, -1.0), (0.0, 0.0, 0.0)],
    "trigonal planar": [(1.0, 0.0, 0.0), (-0.5, 0.0, 0.8660), (-0.5, 0.0, -0.8660), (0.0, 0.0, 0.0)],
    "trigonal bipyramid": [(0.0, 1.0, 0.0), (0.0, -1.0, 0.0), (1.0, 0.0, 0.0), (-0.5, 0.0, 0.8660), (-0.5, 0.0, -0.8660),
                           (0.0, 0.0, 0.0)],
    "trigonal prism": [(-0.6547, 0.6547, 0.3779), (0.6547, 0.6547, 0.3779), (0.0, 0.6547, -0.7559),
                       (-0.6547, -0.6547, 0.3779), (0.6547, -0.6547, 0.3779), (0.0, -0.6547, -0.7559), (0.0, 0.0, 0.0)],
}


def enable(**kwargs):
    kwargs = Coordination.validate(kwargs)
    return Coordination(**kwargs)


def _residue_spec(x):
    """``'Molecule/position'`` or ``[molecule, position]``; the position may be ``'*'``
    (parse.Named_spec("molecule", "residue"))."""
    if isinstance(x, str):
        mol, pos = x.split("/")
    else:
        mol, pos = x
    return (mol, pos if str(pos) == "*" else int(pos))


class Coordination(ObjectiveProvider):
    _validate = {"probe": "required", "residues": "required", "radius": None, "atom_types": None, "atom_names": None,
                 "atom_elements": None, "distance": None, "min_atoms": None, "geometry": None,
                 "enforce_all_residues": None, "only_one_ligand_per_residue": None, "prevent_intruders": None,
                 "center_of_mass_correction": None, "distance_correction": None}
    coordination = True

    def __init__(self, probe=None, radius=3.0, atom_types=(), atom_elements=(), atom_names=(), residues=(),
                 geometry="tetrahedral", distance=0, min_atoms=1, prevent_intruders=True, enforce_all_residues=False,
                 only_one_ligand_per_residue=False, center_of_mass_correction=False, distance_correction=False,
                 **kwargs):
        ObjectiveProvider.__init__(self, **kwargs)
        self._probe = _spec(probe)
        self._residues = [_residue_spec(r) for r in residues]
        self.radius = float(radius)
        self.atom_types, self.atom_names = tuple(atom_types or ()), tuple(atom_names or ())
        self.atom_elements = tuple(atom_elements or ())
        self.distance = float(distance)
        self.min_atoms = int(min_atoms)
        # stored and never read by the reference's evaluate (coordination.py:141); the same here
        self.only_one_ligand_per_residue = bool(only_one_ligand_per_residue)
        self.enforce_all_residues = bool(enforce_all_residues)
        self.prevent_intruders = bool(prevent_intruders)
        self.center_of_mass_correction = bool(center_of_mass_correction)
        self.distance_correction = bool(distance_correction)
        self.geometry = np.array(GEOMETRIES[geometry] if isinstance(geometry, str) else geometry, np.float64)
        self.n_vertices = self.geometry.shape[0] - 1
        if not (1 <= self.n_vertices <= 8):
            raise ValueError("geometry with %d vertices (1 to 8 supported)" % self.n_vertices)
        if not any((self.atom_types, self.atom_elements, self.atom_names)):
            raise ValueError("At least one of atom_types, atom_elements, atom_names must be specified")

    def site(self, system):
        """The tables of ``gmm_set_coordination`` for this objective.  ``system``: the job's atoms
        (``base.BatchEvaluator.system``: ``xyz``, ``element``, ``idatm``, ``name``, ``molecule``, ``resseq``,
        ``neighbours``, ``mobile_row``, ``locate``)."""
        n = len(system.xyz)
        wanted = {}
        for mol, pos in self._residues:
            wanted.setdefault(mol, set()).add(pos)
        in_residue = np.array([(m in wanted) and ("*" in wanted[m] or int(r) in wanted[m])
                               for m, r in zip(system.molecule, system.resseq)], bool)
        if self.atom_types:                                # the reference's precedence (coordination.py:244-252)
            valid = np.array([t in self.atom_types for t in system.idatm], bool)
        elif self.atom_elements:
            want = {e.upper() for e in self.atom_elements}
            valid = np.array([e.upper() in want for e in system.element], bool)
        else:
            valid = np.array([nm in self.atom_names for nm in system.name], bool)
        return {"mobile_row": system.mobile_row, "xyz": system.xyz, "candidate": (valid & in_residue).astype(np.uint8),
                "geometry": np.array([molio.IDATM_GEOMETRY.get(t, 3) for t in system.idatm], np.uint8),
                "bond_radius": np.array([molio.BOND_RADIUS.get(e.upper(), 0.8) for e in system.element], np.float64),
                "neighbours": system.neighbours, "metal": system.locate(self._probe), "template": self.geometry,
                "radius": self.radius, "distance": self.distance, "weight": float(self.weight if self.weight is not None else 1.0),
                "min_atoms": self.min_atoms, "prevent_intruders": self.prevent_intruders,
                "enforce_all_residues": self.enforce_all_residues,
                "center_of_mass_correction": self.center_of_mass_correction,
                "distance_correction": self.distance_correction}

    def evaluate(self, individual):
        return self.environment.batch.score_of(individual, self)
