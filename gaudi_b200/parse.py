"""Job description: the subset of gaudi/parse.py:162-343 the hot path needs — attribute-style
access to ``output / ga / similarity / genes / objectives`` with the reference's defaults
(parse.py:254-284).  Accepts a dict or a YAML file; the reference's voluptuous validation of
every key is out of scope."""
from __future__ import annotations

import os
from copy import deepcopy


class Param(dict):
    """dict with attribute access (the reference uses ``Munch``)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)

    def __setattr__(self, k, v):
        self[k] = v


DEFAULTS = {
    "output": {"path": ".", "name": "gaudi", "precision": 3, "compress": True, "history": False,
               "pareto": True, "verbose": True, "check_every": 10, "prompt_on_exception": False},
    "ga": {"population": 10, "generations": 3, "mu": 1, "lambda_": 3, "mut_eta": 5, "mut_pb": 0.5,
           "mut_indpb": 0.75, "cx_eta": 5, "cx_pb": 0.5},
    "similarity": {"module": "gaudi_b200.similarity.rmsd", "args": [["Ligand"], 2.5], "kwargs": {}},
    "genes": [], "objectives": [],
}


# A job written for the reference names its plugins ``gaudi.genes.torsion`` etc.  The ones that have a
# B200 mirror are redirected; anything else of the reference's is refused by name, so that a job
# never silently runs with a gene or objective missing.
B200_MODULES = {
    "gaudi.genes.molecule": "gaudi_b200.genes.molecule",
    "gaudi.genes.torsion": "gaudi_b200.genes.torsion",
    "gaudi.genes.search": "gaudi_b200.genes.search",
    "gaudi.genes.rotamers": "gaudi_b200.genes.rotamers",
    "gaudi.objectives.contacts": "gaudi_b200.objectives.contacts",
    "gaudi.objectives.distance": "gaudi_b200.objectives.distance",
    "gaudi.objectives.angle": "gaudi_b200.objectives.angle",
    "gaudi.objectives.coordination": "gaudi_b200.objectives.coordination",
    "gaudi.similarity.rmsd": "gaudi_b200.similarity.rmsd",
}


def b200_module(name):
    """Module path on the B200 path for a plugin named in a reference job file."""
    if name in B200_MODULES:
        return B200_MODULES[name]
    if name.startswith("gaudi."):
        raise NotImplementedError("%s has no B200 counterpart (available: %s)" % (name, ", ".join(sorted(B200_MODULES))))
    return name


class Settings(Param):
    def __init__(self, source=None, **kw):
        Param.__init__(self)
        data = deepcopy(DEFAULTS)
        raw = {}
        self._path = None
        if isinstance(source, str):
            import yaml
            with open(source) as fh:
                raw = yaml.safe_load(fh) or {}
            self._path = os.path.abspath(source)
        elif source:
            raw = deepcopy(dict(source))
        raw.update(kw)
        for key, val in raw.items():
            if isinstance(val, dict) and isinstance(data.get(key), dict):
                data[key].update(val)
            else:
                data[key] = val
        base = os.path.dirname(self._path) if self._path else os.getcwd()
        for key in ("output", "ga", "similarity"):
            self[key] = Param(data[key])
        self["genes"] = [Param(g) for g in data["genes"]]
        self["objectives"] = [Param(o) for o in data["objectives"]]
        if self._path and not os.path.isabs(self["output"]["path"]):    # parse.py:288 RelPathToInputFile
            self["output"]["path"] = os.path.normpath(os.path.join(base, os.path.expanduser(self["output"]["path"])))
        for g in self["genes"]:          # relative paths resolve against the input file (parse.py:130-138)
            if "path" in g and not os.path.isabs(g["path"]):
                g["path"] = os.path.normpath(os.path.join(base, g["path"]))
        for o in self["objectives"]:
            o.setdefault("weight", -1.0)
        for plugin in self["genes"] + self["objectives"]:
            plugin["module"] = b200_module(plugin["module"])
        self["similarity"]["module"] = b200_module(self["similarity"]["module"])

    @property
    def weights(self):
        return [o.weight for o in self.objectives]

    @property
    def name_objectives(self):
        return [o.name for o in self.objectives]
