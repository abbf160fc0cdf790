"""Objective plugins of the B200 path; protocol of gaudi/objectives/__init__.py:43-111:
``enable(**kwargs)`` with kwargs ``environment, zone, precision, name, weight, module`` plus the
objective's YAML keys; one instance per Environment; ``evaluate(individual) -> float``."""
from __future__ import annotations

from uuid import uuid4


class ObjectiveProvider(object):
    _cache = {}
    _validate = {}

    def __init__(self, environment=None, name=None, weight=None, zone=None, precision=3, **kwargs):
        self.environment = environment
        self.name = name if name is not None else str(uuid4())
        self.weight = weight
        self.zone = zone
        self.precision = int(precision)

    def __ready__(self):
        pass

    def evaluate(self, individual):
        raise NotImplementedError

    @classmethod
    def clear_cache(cls):
        cls._cache.clear()

    @classmethod
    def validate(cls, data):
        allowed = set(cls._validate) | {"environment", "module", "name", "weight", "zone", "precision"}
        unknown = set(data) - allowed
        if unknown:
            raise ValueError("%s: unexpected keys %s" % (cls.__name__, sorted(unknown)))
        missing = {k for k, v in cls._validate.items() if v == "required"} - set(data)
        if missing:
            raise ValueError("%s: missing required keys %s" % (cls.__name__, sorted(missing)))
        return dict(data)
