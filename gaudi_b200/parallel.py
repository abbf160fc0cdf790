"""Multi-GPU plumbing: one process per GPU (``torchrun``), population sharded by individual,
one ``all_gather`` of the fitness block per evaluation.  The data path
has no collective: individuals are independent (gaudi/algorithms.py:99-102 is a pure map).
RMSD: coordinates are all-gathered, each rank computes a block of rows.

Works with the ``nccl`` backend on GPUs and with ``gloo`` on CPU tensors (used by the
world-size-2 tests, where the per-shard compute is supplied by the test)."""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_bounds(n, world):
    """Contiguous, balanced blocks: rank r owns [b[r], b[r+1])."""
    return [(n * r) // world for r in range(world + 1)]


def max_shard(n, world):
    b = shard_bounds(n, world)
    return max(b[r + 1] - b[r] for r in range(world))


def sharded_evaluate(evaluate_shard, alleles, n_out=2, device=None, group=None):
    """Evaluate a population split over the ranks of ``group``.

    ``alleles``: tuple of arrays with the population on axis 0 (every rank holds all of them —
    they are tiny — or at least its own slice); ``evaluate_shard(*slices) -> (p, n_out)`` tensor
    or array for this rank's slice.  Returns the full ``(P, n_out)`` float64 tensor on every
    rank (rank 0 is the one that runs NSGA-II selection)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    P = int(alleles[0].shape[0])
    b = shard_bounds(P, world)
    lo, hi = b[rank], b[rank + 1]
    part = evaluate_shard(*[a[lo:hi] if a is not None else None for a in alleles])
    part = torch.as_tensor(part, dtype=torch.float64)
    if device is not None:
        part = part.to(device)
    if world == 1:
        return part
    m = max_shard(P, world)
    padded = torch.zeros((m, n_out), dtype=torch.float64, device=part.device)
    padded[:hi - lo] = part
    gathered = torch.empty((world * m, n_out), dtype=torch.float64, device=part.device)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    out = torch.cat([gathered[r * m:r * m + (b[r + 1] - b[r])] for r in range(world)])
    return out


def sharded_rmsd_bits(rmsd_rows, coords, group=None, counts=None):
    """Row-block sharded similarity bitmask.  ``coords``: this rank's (n_r, A, 3) float32 block
    of conformers; ``rmsd_rows(all_coords, row0, row1)`` returns the (row1-row0, ceil(N/32)) int32
    words for those rows.  ``counts``: conformers per rank when the caller knows them (blocks that
    follow ``shard_bounds``); without it the ranks exchange their block sizes first, which costs a
    host round trip per rank.  Returns the full (N, words) tensor on every rank."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1:
        return rmsd_rows(coords, 0, coords.shape[0])
    if counts is None:
        sizes = torch.zeros(world, dtype=torch.int64, device=coords.device)
        dist.all_gather_into_tensor(sizes, torch.tensor([coords.shape[0]], dtype=torch.int64, device=coords.device),
                                    group=group)
        counts = sizes.tolist()                     # one device-to-host copy
    counts = [int(c) for c in counts]
    m = max(counts)
    even = min(counts) == m                         # equal blocks: gather in place, no padding, no re-packing
    if even:
        padded = coords.contiguous()
    else:
        padded = torch.zeros((m,) + tuple(coords.shape[1:]), dtype=coords.dtype, device=coords.device)
        padded[:coords.shape[0]] = coords
    allc = torch.empty((world * m,) + tuple(coords.shape[1:]), dtype=coords.dtype, device=coords.device)
    dist.all_gather_into_tensor(allc, padded, group=group)
    if not even:
        allc = torch.cat([allc[r * m:r * m + counts[r]] for r in range(world)])
    n = allc.shape[0]
    row0 = sum(counts[:rank])
    mine = rmsd_rows(allc, row0, row0 + counts[rank])
    words = (n + 31) // 32
    if even:
        pad = mine
    else:
        pad = torch.zeros((m, words), dtype=mine.dtype, device=mine.device)
        pad[:counts[rank]] = mine
    allb = torch.empty((world * m, words), dtype=mine.dtype, device=mine.device)
    dist.all_gather_into_tensor(allb, pad, group=group)
    if even:
        return allb
    return torch.cat([allb[r * m:r * m + counts[r]] for r in range(world)])
