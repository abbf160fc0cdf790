"""Gene plugins of the B200 path.  Same protocol as the reference's ``GeneProvider``
(gaudi/genes/__init__.py:44-139): a module-level ``enable(**kwargs)`` returning an instance
that owns an ``allele`` and implements ``express / unexpress / mate / mutate / write``; one
instance per individual, created through ``gaudi_b200.plugin.load_plugins`` with the kwargs
``parent, name, cx_eta, mut_eta, mut_indpb`` plus the gene's YAML keys.

Difference to the reference: ``express()`` of a single gene does not touch any geometry
here.  Expression happens for the whole population at once in
``gaudi_b200.base.BatchEvaluator`` (one ``gmm_evaluate`` call); per-gene ``express`` only
marks the individual as expressed so that code written against the reference protocol
(``with expressed(ind): ...``) keeps working.
"""
from __future__ import annotations

import os
import pprint
from uuid import uuid4

pp = pprint.PrettyPrinter(4)


class GeneProvider(object):
    _cache = {}
    _validate = {}

    def __init__(self, parent=None, name=None, cx_eta=5.0, mut_eta=5.0, mut_indpb=0.75, **kwargs):
        self.parent = parent
        self.name = name if name is not None else str(uuid4())
        self.cxeta = float(cx_eta)
        self.mteta = float(mut_eta)
        self.indpb = float(mut_indpb)
        self.allele = None

    def __ready__(self):
        pass

    def __expression_hooks__(self):
        pass

    def express(self):
        pass

    def unexpress(self):
        pass

    def mutate(self, indpb):
        raise NotImplementedError

    def mate(self, gene):
        raise NotImplementedError

    @classmethod
    def validate(cls, data):
        """Light-weight stand-in for the voluptuous schemas (gaudi/parse.py:375-380): unknown
        keys are rejected, required ones demanded."""
        allowed = set(cls._validate) | {"parent", "name", "module", "cx_eta", "mut_eta", "mut_indpb"}
        required = {k for k, v in cls._validate.items() if v == "required"}
        unknown = set(data) - allowed
        if unknown:
            raise ValueError("%s: unexpected keys %s" % (cls.__name__, sorted(unknown)))
        missing = required - set(data)
        if missing:
            raise ValueError("%s: missing required keys %s" % (cls.__name__, sorted(missing)))
        return dict(data)

    def write(self, path, name, *args, **kwargs):
        fullname = os.path.join(path, "{}_{}.txt".format(name, self.name))
        with open(fullname, "w") as f:
            f.write(pp.pformat(self.allele))
        return fullname

    @classmethod
    def clear_cache(cls):
        cls._cache.clear()
