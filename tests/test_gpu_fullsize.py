"""BASELINE.json configs[3] and configs[4] at their stated sizes, on the GPU through the C ABI.

configs[3]: hall-of-fame pairwise RMSD of 16,384 conformers x 60 atoms, row-block sharded.
configs[4]: Rotamers gene on 12 flexible side chains + Contacts, population 32,768.
(configs[0]-[2] at full size live in test_gpu_plugins.py / test_gpu_parity.py.)"""
import os
import random
import sys

import numpy as np
import pytest

import golden_util as G
import oracle_adapter as OA
from gaudi_b200 import model as M
from gaudi_b200 import synth

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCORE_RTOL = 1e-4


@pytest.fixture(scope="module")
def torch_cuda(built_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch


def test_configs3_similarity_16384_conformers(torch_cuda):
    """gaudi/similarity.py:36-75 over every pair of 16,384 conformers of 60 atoms (134 M unique
    pairs), threshold 2.5 A as gaudi/parse.py:279 sets it:
      - the tensor-core bitmask equals the FP32 direct-difference kernel's word for word;
      - 4,096 sampled pairs equal the fp64 oracle (every sampled pair further than 1e-4 A from the
        threshold; the RMSD values of the same pairs within 1e-5 A);
      - eight row blocks (what eight GPUs compute) are the rows of the full matrix;
      - the partner lists compacted per row block concatenate to the full matrix's lists;
      - symmetric, diagonal set."""
    torch = torch_cuda
    from gaudi_b200.engine import Engine, bits_to_bool
    from gaudi_b200 import parallel
    n, a, thr = 16384, 60, 2.5
    lig = synth.make_ligand(12, 2, seed=1)
    conf = synth.make_conformers(n, a, seed=4).astype(np.float32)
    conf_d = torch.as_tensor(conf).cuda()
    eng = Engine(lig, M.empty_receptor(), M.ContactParams(only_internal=True), device=0)
    simt = Engine(lig, M.empty_receptor(), M.ContactParams(only_internal=True, rmsd_simt=True), device=0)
    eng.reserve_rmsd(n, a)
    _, bits = eng.rmsd_matrix(conf_d, threshold=thr, values=False)
    _, bits_simt = simt.rmsd_matrix(conf_d, threshold=thr, values=False)
    assert torch.equal(bits, bits_simt)
    # sampled pairs against the fp64 oracle (the formula of similarity.py:73-75)
    rng = np.random.default_rng(44)
    ii, jj = rng.integers(0, n, 4096), rng.integers(0, n, 4096)
    # a third of the sample from the interesting region: partners the bitmask calls similar
    ptr, idx = eng.similar_lists(bits, n)
    rows_with = np.flatnonzero(np.diff(ptr) > 0)
    pick = rng.choice(rows_with, 1365)
    ii[:1365] = pick
    jj[:1365] = idx[ptr[pick] + (rng.integers(0, 1 << 30, 1365) % np.diff(ptr)[pick])]
    c64 = conf.astype(np.float64)
    ref = np.sqrt(((c64[ii] - c64[jj]) ** 2).sum(axis=(1, 2)) / a)
    words = bits.cpu().numpy().view(np.uint32)
    got = ((words[ii, jj >> 5] >> (jj & 31).astype(np.uint32)) & 1).astype(bool)
    clear = np.abs(ref - thr) > 1e-4
    assert clear.sum() > 4000 and got[:1365].all() and (~got[1365:]).sum() > 2000
    assert np.array_equal(got[clear], (ref <= thr)[clear])
    # values of one row block against the oracle formula, 1e-5 A
    vals, _ = eng.rmsd_matrix(conf_d, row0=4096, row1=4096 + 64, threshold=thr, bits=False)
    vals = vals.cpu().numpy()
    for r in range(64):
        ref_row = np.sqrt(((c64[4096 + r][None] - c64) ** 2).sum(axis=(1, 2)) / a)
        assert np.abs(vals[r] - ref_row).max() < 1e-5, r
    # row blocks of 1/8 each == the rows of the full matrix; so are their partner lists
    b = parallel.shard_bounds(n, 8)
    all_ptr, all_idx = [np.zeros(1, np.int64)], []
    for r in range(8):
        _, blk = eng.rmsd_matrix(conf_d, row0=b[r], row1=b[r + 1], threshold=thr, values=False)
        assert torch.equal(blk, bits[b[r]:b[r + 1]]), r
        p, i = eng.similar_lists(blk, n, row0=b[r])
        all_ptr.append(p[1:] + all_ptr[-1][-1]); all_idx.append(i)
    assert np.array_equal(np.concatenate(all_ptr), ptr) and np.array_equal(np.concatenate(all_idx), idx)
    # symmetric with the diagonal set: partner lists are mutual
    assert bool(((words[np.arange(n), np.arange(n) >> 5] >> (np.arange(n) & 31).astype(np.uint32)) & 1).all())
    src = np.repeat(np.arange(n), np.diff(ptr))
    fwd = set(zip(src[:20000].tolist(), idx[:20000].tolist()))
    back = ((words[idx[:20000], src[:20000] >> 5] >> (src[:20000] & 31).astype(np.uint32)) & 1).astype(bool)
    assert back.all() and len(fwd) == 20000
    eng.close(); simt.close()


def test_configs4_rotamers_12_residues_pop_32768(torch_cuda):
    """Rotamers.express (gaudi/genes/rotamers.py:141-152,186-210) on 12 flexible side chains of the
    5er1 complex + Torsion + Search, Contacts with the CB zone seeds of
    gaudi/objectives/contacts.py:195-207, population 32,768: EVERY individual against the oracle's C
    restatement (coordinates 1e-4 A, scores 1e-4 relative, pair counts equal); individuals that
    differ must be explained by pairs within 1e-4 A of a cutoff, which the numpy oracle adapter
    quantifies for each of them."""
    torch = torch_cuda
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import c_oracle
    from gaudi_b200 import base
    cfg, residues = G.configs4_job(12, seed=5)
    random.seed(5)
    template = base.MolecularIndividual(cfg)
    env = base.Environment(cfg, device=0)
    batch = env.attach(template)
    eng = batch.engines[0][0]
    assert batch.n_rot_res == 12 and len(residues) == 12
    assert eng.mobile.n_atoms > 150 and len(eng.receptor.ms_excl_idx) > 0
    pop = 32768
    tors, xf, rot = G.configs4_alleles(batch, template, pop, seed=5)
    s, c, x = eng.evaluate(tors, xf, rot, want_xyz=True)
    s = s.cpu().numpy(); c = c.cpu().numpy(); x = x.cpu().numpy().astype(np.float64) + eng.centre
    ref_s, ref_c, ref_x = c_oracle.COracle(eng.mobile, eng.receptor, eng.params).evaluate(tors, xf, rot, want_xyz=True)
    assert np.abs(x - ref_x).max() < 1e-4
    rel = np.abs(s - ref_s) / np.maximum(1.0, np.abs(ref_s))
    bad = np.flatnonzero((rel > SCORE_RTOL).any(axis=1) | (c != ref_c).any(axis=1))
    print("configs[4]: %d of %d individuals differ from the C oracle beyond 1e-4 relative / pair counts" % (len(bad), pop))
    assert len(bad) < 0.02 * pop, len(bad)
    for p in bad[:48]:
        res = OA.contacts(eng.mobile, eng.receptor, eng.params, ref_x[p])
        assert res.borderline, (p, s[p], ref_s[p])
        assert abs(s[p, 0] - res.clash) <= res.clash_slack + SCORE_RTOL * max(1.0, abs(res.clash)), (p, s[p], res.clash)
        assert abs(s[p, 1] - res.hydro) <= res.hydro_slack + SCORE_RTOL * max(1.0, abs(res.hydro)), (p, s[p], res.hydro)
    # every ligand atom of these poses left its template position (Search + Torsion acted)
    moved = np.abs(x[:64] - eng.mobile.xyz0[None]).max(axis=(0, 2))
    assert (moved[batch.lig_slice] > 1.0).all()
    # sharding invariance at this size: four shards give the same rows
    parts = [eng.evaluate(tors[k * 8192:(k + 1) * 8192], xf[k * 8192:(k + 1) * 8192], rot[k * 8192:(k + 1) * 8192])[0]
             for k in range(4)]
    assert np.array_equal(torch.cat(parts).cpu().numpy(), s)
    batch.close()
