"""Contacts objective (gaudi/objectives/contacts.py): same parameters and defaults
(contacts.py:101-121).  ``evaluate(individual)`` answers from the batch the individual was
scored in (``BatchEvaluator``); scoring one individual on its own goes through the same GPU
path with a population of one."""
from __future__ import annotations

from ..model import ContactParams
from . import ObjectiveProvider


def enable(**kwargs):
    kwargs = Contacts.validate(kwargs)
    return Contacts(**kwargs)


class Contacts(ObjectiveProvider):
    _validate = {"probes": "required", "radius": None, "which": None, "clash_threshold": None,
                 "hydrophobic_threshold": None, "cutoff": None, "hydrophobic_elements": None,
                 "bond_separation": None, "same_residue": None, "only_internal": None}

    def __init__(self, probes=None, radius=5.0, which="hydrophobic", clash_threshold=0.6,
                 hydrophobic_threshold=-0.4, cutoff=0.0, hydrophobic_elements=("C", "S"), bond_separation=4,
                 same_residue=True, only_internal=False, **kwargs):
        ObjectiveProvider.__init__(self, **kwargs)
        if which not in ("hydrophobic", "clashes"):
            raise ValueError("Contacts: which must be 'hydrophobic' or 'clashes' (contacts.py:91)")
        if float(radius) < 0:
            raise ValueError("Contacts: radius must be >= 0 (contacts.py:90)")
        if int(bond_separation) < 2:
            raise ValueError("Contacts: bond_separation must be >= 2 (contacts.py:96)")
        self.which = which
        self.radius = float(radius)
        self.clash_threshold = float(clash_threshold)
        self.hydrophobic_threshold = float(hydrophobic_threshold)
        self.cutoff = float(cutoff)
        self.hydrophobic_elements = tuple(hydrophobic_elements)
        self.bond_separation = int(bond_separation)
        self.same_residue = bool(same_residue)
        self.only_internal = bool(only_internal)
        self._probes = list(probes)
        self.threshold = hydrophobic_threshold if which == "hydrophobic" else clash_threshold

    @property
    def column(self):
        """Column of the engine's (clashes, hydrophobic) output this objective reads."""
        return 0 if self.which == "clashes" else 1

    def model_key(self):
        """Objectives with the same key share one GPU context (both scores come out of one pass)."""
        return (tuple(self._probes), self.radius, self.clash_threshold, self.hydrophobic_threshold,
                tuple(sorted(self.hydrophobic_elements)), self.bond_separation, self.only_internal, self.same_residue)

    def penalty_key(self):
        """Cutoff penalty of evaluate_clashes (contacts.py:132-133); irrelevant for hydrophobic."""
        if self.which == "clashes" and self.cutoff:
            return (self.cutoff, self.weight)
        return None

    def params(self):
        return ContactParams(radius=self.radius, clash_threshold=self.clash_threshold,
                             hydrophobic_threshold=self.hydrophobic_threshold, cutoff=self.cutoff,
                             weight=self.weight if self.weight is not None else 1.0,
                             bond_separation=self.bond_separation, only_internal=self.only_internal,
                             same_residue=self.same_residue)

    def evaluate(self, individual):
        return self.environment.batch.score_of(individual, self)
