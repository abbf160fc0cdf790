"""Angle objective (gaudi/objectives/angle.py): deviation of an angle (3 atoms) or dihedral
(4 atoms) from a threshold, or |sin| for ``threshold: planar``; evaluated by ``gmm_geometry``."""
from __future__ import annotations

from . import ObjectiveProvider
from .distance import _spec


def enable(**kwargs):
    kwargs = Angle.validate(kwargs)
    return Angle(**kwargs)


class Angle(ObjectiveProvider):
    _validate = {"probes": "required", "threshold": "required"}
    geometry = True

    def __init__(self, threshold=None, probes=None, **kwargs):
        ObjectiveProvider.__init__(self, **kwargs)
        self.threshold = threshold if threshold == "planar" else float(threshold)
        self._probes = [_spec(p) for p in probes]
        if len(self._probes) not in (3, 4):
            raise ValueError("Angle: probes must list 3 (angle) or 4 (dihedral) atoms")

    def terms(self, resolve, covalent):
        return [{"kind": 1 if len(self._probes) == 3 else 2, "points": [resolve(p) for p in self._probes],
                 "threshold": self.threshold, "weight": self.weight}]

    def evaluate(self, individual):
        return self.environment.batch.score_of(individual, self)
