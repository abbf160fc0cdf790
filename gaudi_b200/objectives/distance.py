"""Distance objective (gaudi/objectives/distance.py): mean absolute deviation of probe-target
distances from a threshold, evaluated for the whole population by ``gmm_geometry``."""
from __future__ import annotations

from . import ObjectiveProvider


def enable(**kwargs):
    kwargs = Distance.validate(kwargs)
    return Distance(**kwargs)


def _is_atom(x):
    return (isinstance(x, str) and "/" in x) or \
        (isinstance(x, (list, tuple)) and len(x) == 2 and isinstance(x[0], str))


def _spec(x):
    """``'Molecule/serial'`` or ``[molecule, serial]`` -> ``(molecule, serial)`` (parse.Named_spec)."""
    if isinstance(x, str) and "/" in x:
        mol, serial = x.split("/")
        return (mol, int(serial))
    if _is_atom(x):
        return (x[0], int(x[1]))
    return x


class Distance(ObjectiveProvider):
    _validate = {"probes": "required", "target": "required", "threshold": "required", "tolerance": None,
                 "center_of_mass": None}
    geometry = True

    def __init__(self, threshold=None, tolerance=None, target=None, probes=None, center_of_mass=False, **kwargs):
        ObjectiveProvider.__init__(self, **kwargs)
        self.center_of_mass = bool(center_of_mass)
        self.threshold = threshold if threshold == "covalent" else float(threshold)
        self.tolerance = None if tolerance is None else float(tolerance)
        self._probes = [_spec(p) for p in ([probes] if _is_atom(probes) else probes)]
        self._target = _spec(target)

    def terms(self, resolve, covalent, mass=None):
        """``resolve((mol, serial))`` -> mobile row (int) or world xyz of a static atom;
        ``covalent(a, b)`` -> bond length; ``mass(atom)`` -> atomic mass (``center_of_mass``)."""
        target = resolve(self._target) if isinstance(self._target[0], str) else [float(v) for v in self._target]
        out = []
        if self.center_of_mass:
            # distance.py:131-148: distance of the target to the mass-weighted centre of the probes
            return [{"kind": 3, "points": [resolve(p), target], "threshold": mass(p), "tolerance": None,
                     "weight": self.weight} for p in self._probes]
        for p in self._probes:
            thr = covalent(p, self._target) if self.threshold == "covalent" else self.threshold
            out.append({"kind": 0, "points": [resolve(p), target], "threshold": thr, "tolerance": self.tolerance,
                        "weight": self.weight})
        return out

    def evaluate(self, individual):
        return self.environment.batch.score_of(individual, self)
