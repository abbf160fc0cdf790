"""``similarity.rmsd`` with the reference's signature (gaudi/similarity.py:36-70):
``rmsd(ind1, ind2, subjects, threshold) -> bool``, no superposition, ``False`` when the atom
counts differ.  Inside the GA loop the answer comes from the bitmask that one
``gmm_rmsd_matrix`` call computed for the whole generation
(``BatchEvaluator.similarity_matrix``); a pair that is not in that batch is computed on the GPU
on its own (a 2-conformer matrix).  There is no CPU path."""
from __future__ import annotations

import numpy as np

_engines_for_pairs = {}


def _pair_engine(batch=None, device=None):
    """The context stand-alone RMSD calls run on: the job's own (first) engine when there is one,
    else a context with a dummy one-atom model on ``device`` (default: the current CUDA device, so
    that under one-process-per-GPU launches no rank touches another rank's GPU)."""
    if batch is not None and batch.engines:
        return batch.engines[0][0]
    if device is None:
        import torch
        device = torch.cuda.current_device()
    device = int(device)
    if device not in _engines_for_pairs:
        from . import model as M
        from .engine import Engine
        mob = M.build_mobile(np.zeros((1, 3)), np.ones(1), np.array([M.F_SEED], np.uint8), np.zeros((0, 2)), [])
        _engines_for_pairs[device] = Engine(mob, M.empty_receptor(), M.ContactParams(only_internal=True), device=device)
    return _engines_for_pairs[device]


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
        _, bits = _pair_engine(batch, getattr(env, "device", None)).rmsd_matrix(conf, threshold=threshold, values=False)
        if not bits_to_bool(bits, 2)[0, 1]:
            return False
    return True
