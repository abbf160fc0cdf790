"""``similarity.rmsd`` with the reference's signature (gaudi/similarity.py:36-70):
``rmsd(ind1, ind2, subjects, threshold) -> bool``, no superposition, ``False`` when the atom
counts differ.  Inside the GA loop the answer comes from the bitmask that one
``gmm_rmsd_matrix`` call computed for the whole generation
(``BatchEvaluator.similarity_matrix``); a pair that is not in that batch is computed on the GPU
on its own (a 2-conformer matrix).  There is no CPU path."""
from __future__ import annotations

import numpy as np

_engine_for_pairs = None


def _pair_engine():
    """A context with a dummy one-atom model, only used for stand-alone RMSD calls."""
    global _engine_for_pairs
    if _engine_for_pairs is None:
        from . import model as M
        from .engine import Engine
        mob = M.build_mobile(np.zeros((1, 3)), np.ones(1), np.array([M.F_SEED], np.uint8), np.zeros((0, 2)), [])
        _engine_for_pairs = Engine(mob, M.empty_receptor(), M.ContactParams(only_internal=True))
    return _engine_for_pairs


def rmsd(ind1, ind2, subjects, threshold, *args, **kwargs):
    batch = None
    env = getattr(getattr(ind1, "cfg", None), "_environment", None)
    if env is not None:
        batch = env.batch
    if batch is not None:
        hit = batch.lookup_similarity(ind1, ind2, subjects, threshold)
        if hit is not None:
            return hit
    from .engine import bits_to_bool
    for s in subjects:
        c1 = ind1.find_molecule(s)._expressed_coordinates
        c2 = ind2.find_molecule(s)._expressed_coordinates
        if c1.shape[0] != c2.shape[0]:
            return False                       # similarity.py:65-66
        centre = c1.mean(axis=0)
        conf = np.stack([c1 - centre, c2 - centre]).astype(np.float32)
        _, bits = _pair_engine().rmsd_matrix(conf, threshold=threshold, values=False)
        if not bits_to_bool(bits, 2)[0, 1]:
            return False
    return True
