"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same
seeded inputs.  Tolerances are the ones BASELINE.json's north_star states:
coordinates 1e-4 A, Contacts scores 1e-4 relative, identical pair sets except pairs within
1e-4 A of a cutoff, RMSD 1e-5 A."""
import numpy as np
import pytest

import oracle_adapter as OA
from gaudi_b200 import model as M
from gaudi_b200 import synth

pytestmark = pytest.mark.gpu

XYZ_TOL = 1e-4      # A
SCORE_RTOL = 1e-4   # relative
RMSD_TOL = 1e-5     # A


@pytest.fixture(scope="module")
def torch_cuda(built_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch


@pytest.fixture(scope="module")
def small_system(torch_cuda):
    from gaudi_b200.engine import Engine
    params = M.ContactParams()
    lig = synth.make_ligand(40, 10, seed=1)
    rec = synth.make_receptor(5000, seed=2, params=params)
    eng = Engine(lig, rec, params, device=0)
    return lig, rec, params, eng


def score_close(got, want, slack):
    err = abs(got - want)
    return err <= slack + SCORE_RTOL * max(1.0, abs(want))


def test_express_matches_oracle(small_system, torch_cuda):
    lig, rec, params, eng = small_system
    tors, xf = synth.make_alleles(4096, lig.n_allele_torsions, seed=3)
    xyz = eng.express(tors, xf).cpu().numpy().astype(np.float64) + eng.centre
    worst = 0.0
    for p in range(0, 4096, 16):
        ref = OA.express(lig, tors[p], xf[p])
        worst = max(worst, float(np.abs(ref - xyz[p]).max()))
    assert worst < XYZ_TOL, worst
    # size-independent property on the whole population: bond lengths are preserved
    b = lig.bonds
    d0 = np.linalg.norm(lig.xyz0[b[:, 0]] - lig.xyz0[b[:, 1]], axis=1)
    d = np.linalg.norm(xyz[:, b[:, 0]] - xyz[:, b[:, 1]], axis=2)
    assert np.abs(d - d0).max() < 2e-5
    # ... and the Search target atom lands on the translation of the allele (search.py:190-214)
    assert np.abs(xyz[:, lig.origin_atom] - xf.reshape(-1, 3, 4)[:, :, 3]).max() < 1e-5


def test_express_tree_form_and_sequential_agree(small_system, torch_cuda):
    """The kinematic-tree express kernel (template axes, all bond maps built at once) against the
    kernel that applies the rotations one after another, as the reference does; and a model whose
    moving sets do NOT form a tree, which must be detected and expressed sequentially."""
    import copy
    from gaudi_b200.engine import Engine
    lig, rec, params, eng = small_system
    tors, xf = synth.make_alleles(777, lig.n_allele_torsions, seed=31)
    seq = Engine(lig, rec, M.ContactParams(sequential_express=True), device=0)
    a = eng.express(tors, xf)
    b = seq.express(tors, xf)
    assert eng.express_is_tree() and not seq.express_is_tree()
    assert float((a - b).abs().max()) <= 2e-6          # same fp64 maths up to rounding, fp32 output
    # overlapping (not nested) moving sets: bond 1 also drags half of bond 0's moving side along
    odd = copy.deepcopy(lig)
    odd.tors_mask = lig.tors_mask.copy()
    m0 = lig.moving_indices(0)
    extra = [int(i) for i in m0[: max(1, len(m0) // 2)] if i not in (int(lig.tors_fixed[1]), int(lig.tors_moving[1]))]
    for i in extra:
        odd.tors_mask[1, i >> 6] |= np.uint64(1) << np.uint64(i & 63)
    e_odd = Engine(odd, rec, params, device=0)
    assert not e_odd.express_is_tree()
    got = e_odd.express(tors[:64], xf[:64]).cpu().numpy().astype(np.float64) + e_odd.centre
    for p in range(0, 64, 7):
        assert np.abs(OA.express(odd, tors[p], xf[p]) - got[p]).max() < XYZ_TOL


@pytest.mark.parametrize("n_atoms, n_tors, tree", [(120, 40, True), (200, 70, False)])
def test_express_many_torsions(torch_cuda, n_atoms, n_tors, tree):
    """More than 32 bonds (two rounds of per-lane bond maps, 64-bit bond masks) and more than 64
    (past what the tree form holds: sequential kernel)."""
    from gaudi_b200.engine import Engine
    lig = synth.make_ligand(n_atoms, n_tors, seed=9)
    assert lig.n_allele_torsions == n_tors
    eng = Engine(lig, M.empty_receptor(), M.ContactParams(only_internal=True), device=0)
    assert eng.express_is_tree() == tree
    tors, xf = synth.make_alleles(40, n_tors, seed=4)
    tors[3] = 0.0                                    # an individual with every bond at rest
    tors[5, ::2] = 0.0                               # and one with every other bond at rest
    xyz = eng.express(tors, xf).cpu().numpy().astype(np.float64) + eng.centre
    for p in range(0, 40, 3):
        assert np.abs(OA.express(lig, tors[p], xf[p]) - xyz[p]).max() < XYZ_TOL


def test_express_identity(small_system):
    lig, rec, params, eng = small_system
    tors = np.zeros((3, lig.n_allele_torsions))
    xyz = eng.express(tors, None).cpu().numpy().astype(np.float64) + eng.centre
    assert np.abs(xyz - lig.xyz0).max() < 1e-5


@pytest.mark.parametrize("cull", [True, False])
def test_contacts_match_oracle(small_system, torch_cuda, cull):
    from gaudi_b200.engine import Engine
    lig, rec, params, eng = small_system
    if not cull:
        params = M.ContactParams(cull=False)
        eng = Engine(lig, rec, params, device=0)
    n = 96
    tors, xf = synth.make_alleles(n, lig.n_allele_torsions, seed=11)
    scores, counts, xyz = eng.evaluate(tors, xf, want_xyz=True)
    scores = scores.cpu().numpy(); counts = counts.cpu().numpy()
    xyz = xyz.cpu().numpy().astype(np.float64) + eng.centre
    clean = 0
    for p in range(n):
        res = OA.contacts(lig, rec, params, xyz[p])
        assert score_close(scores[p, 0], res.clash, res.clash_slack), (p, scores[p], res.clash, res.clash_slack)
        assert score_close(scores[p, 1], res.hydro, res.hydro_slack), (p, scores[p], res.hydro, res.hydro_slack)
        if not res.borderline:
            clean += 1
            assert counts[p, 0] == res.n_clash and counts[p, 1] == res.n_hydro, (p, counts[p], res.n_clash, res.n_hydro)
    assert clean > n // 2


def test_pair_sets_identical(small_system):
    lig, rec, params, eng = small_system
    tors, xf = synth.make_alleles(6, lig.n_allele_torsions, seed=5)
    xyz32 = eng.express(tors, xf)
    xyz = xyz32.cpu().numpy().astype(np.float64) + eng.centre
    m = lig.n_atoms
    radius = np.concatenate([lig.radius, rec.radius]).astype(np.float64)
    seeds = np.flatnonzero(np.concatenate([lig.flags, rec.flags]) & M.F_SEED)

    def near_a_cutoff(coords, key, ov):
        """The only reasons north_star accepts for a pair to be in one set and not the other: an
        overlap within 1e-4 A of a DetectClash threshold, a distance within 1e-4 A of the
        assumedMaxVdw search bound, or an atom within 1e-4 A of the zone radius (2e-4 allowed:
        fp32 coordinates on one side, fp64 on the other)."""
        a, b = key
        d = float(np.linalg.norm(coords[a] - coords[b]))
        if min(abs(ov - params.clash_threshold), abs(ov - params.hydrophobic_threshold)) < 2e-4:
            return True
        for thr in (params.clash_threshold, params.hydrophobic_threshold):
            if abs(d - (max(radius[a], radius[b]) + params.assumed_max_vdw - thr)) < 2e-4:
                return True
        for atom in (a, b):
            dmin = np.linalg.norm(coords[seeds] - coords[atom], axis=1).min()
            if abs(dmin - params.radius) < 2e-4:
                return True
        return False

    for p in range(6):
        res = OA.contacts(lig, rec, params, xyz[p])
        got = eng.contact_pairs(xyz32[p])
        coords = np.concatenate([xyz[p], rec.xyz])
        got_c = {(a, b): (ov, v) for a, b, k, ov, v in got if k == 0}
        got_h = {(a, b): (ov, v) for a, b, k, ov, v in got if k == 1}
        want_c = {(a, b): (ov, v) for a, b, ov, v in res.neg}
        want_h = {(a, b): (ov, v) for a, b, ov, v in res.pos}
        for got_d, want_d in ((got_c, want_c), (got_h, want_h)):
            for key in set(got_d) ^ set(want_d):
                ov = (got_d.get(key) or want_d.get(key))[0]
                # only pairs within 1e-4 A of a cutoff may differ
                assert res.borderline and near_a_cutoff(coords, key, ov), (p, key, ov)
            if not res.borderline:
                assert set(got_d) == set(want_d)
            for key in set(got_d) & set(want_d):
                assert abs(got_d[key][0] - want_d[key][0]) < 1e-4
                assert abs(got_d[key][1] - abs(want_d[key][1]) if got_d is got_c else got_d[key][1] - want_d[key][1]) \
                    <= SCORE_RTOL * max(1.0, abs(want_d[key][1])) + 1e-4 * 40


def test_contacts_invariances(small_system, torch_cuda):
    """Size-independent properties at a BASELINE-size population: culled == dense bit for bit,
    and any split of the population gives the same rows (sharding invariance)."""
    from gaudi_b200.engine import Engine
    lig, rec, params, eng = small_system
    dense = Engine(lig, rec, M.ContactParams(cull=False), device=0)
    tors, xf = synth.make_alleles(4096, lig.n_allele_torsions, seed=3)
    eng.set_profiling(True); dense.set_profiling(True)        # the pair-test tally is kept while profiling
    s1, c1, _ = eng.evaluate(tors, xf)
    s2, c2, _ = dense.evaluate(tors, xf)
    assert torch_cuda.equal(s1, s2) and torch_cuda.equal(c1, c2)
    assert 0 < eng.last_pair_tests() < dense.last_pair_tests()
    eng.set_profiling(False)
    parts = [eng.evaluate(tors[a:b], xf[a:b])[0] for a, b in ((0, 1000), (1000, 1003), (1003, 4096))]
    assert torch_cuda.equal(torch_cuda.cat(parts), s1)


@pytest.mark.parametrize("pop", [4096, 100, 1])
def test_fused_equals_tiled(small_system, torch_cuda, pop):
    """The one-kernel Contacts path (warp or CTA per individual, expanded-form sweep with exact
    re-decision near the zone radius) and the two-kernel tiled path (direct differences) must
    agree bit for bit: both sum the same per-pair fixed-point values."""
    from gaudi_b200.engine import Engine
    lig, rec, params, eng = small_system
    tors, xf = synth.make_alleles(pop, lig.n_allele_torsions, seed=17)
    s1, c1, x1 = eng.evaluate(tors, xf, want_xyz=True)
    for kw in (dict(tiled=True), dict(tiled=True, cull=False), dict(cull=False)):
        other = Engine(lig, rec, M.ContactParams(**kw), device=0)
        s2, c2 = other.contacts(x1)
        assert torch_cuda.equal(s1, s2) and torch_cuda.equal(c1, c2), kw
        other.close()


def test_evaluate_host_equals_device(small_system):
    lig, rec, params, eng = small_system
    tors, xf = synth.make_alleles(777, lig.n_allele_torsions, seed=9)
    s_dev, c_dev, x_dev = eng.evaluate(tors, xf, want_xyz=True)
    s_h, c_h, x_h = eng.evaluate_host(tors, xf, want_xyz=True)
    assert np.array_equal(s_dev.cpu().numpy(), s_h)
    assert np.array_equal(c_dev.cpu().numpy(), c_h)
    assert np.array_equal(x_dev.cpu().numpy(), x_h)


def test_edge_cases(small_system, torch_cuda):
    from gaudi_b200.engine import Engine
    from gaudi_b200 import _lib
    lig, rec, params, eng = small_system
    # empty population
    s, c, _ = eng.evaluate(np.zeros((0, lig.n_allele_torsions)), np.zeros((0, 12)))
    assert s.shape == (0, 2)
    # single individual, ragged (not a multiple of any tile)
    tors, xf = synth.make_alleles(1, lig.n_allele_torsions, seed=2)
    s, c, x = eng.evaluate(tors, xf, want_xyz=True)
    res = OA.contacts(lig, rec, params, x[0].cpu().numpy().astype(np.float64) + eng.centre)
    assert score_close(float(s[0, 0]), res.clash, res.clash_slack)
    # no receptor at all: only intra-ligand pairs
    e2 = Engine(lig, M.empty_receptor(), params, device=0)
    s2, c2, x2 = e2.evaluate(tors, xf, want_xyz=True)
    res2 = OA.contacts(lig, M.empty_receptor(), params, x2[0].cpu().numpy().astype(np.float64))
    assert score_close(float(s2[0, 0]), res2.clash, res2.clash_slack)
    assert score_close(float(s2[0, 1]), res2.hydro, res2.hydro_slack)
    # only_internal ignores the receptor (contacts.py:199)
    e3 = Engine(lig, rec, M.ContactParams(only_internal=True), device=0)
    s3, _, _ = e3.evaluate(tors, xf)
    assert torch_cuda.equal(s3, s2)
    # bad arguments are reported, not crashed on
    with pytest.raises(_lib.GmmError):
        eng.rmsd_matrix(torch_cuda.zeros((4, 3, 3), device="cuda"), row0=3, row1=9)


def test_cutoff_penalty(small_system):
    from gaudi_b200.engine import Engine
    lig, rec, params, eng = small_system
    pen = Engine(lig, rec, M.ContactParams(cutoff=300.0, weight=-1.0), device=0)
    tors, xf = synth.make_alleles(64, lig.n_allele_torsions, seed=21)
    s0 = eng.evaluate(tors, xf)[0].cpu().numpy()
    s1 = pen.evaluate(tors, xf)[0].cpu().numpy()
    over = s0[:, 0] > 300.0
    assert over.any() and (~over).any()
    assert np.all(s1[over, 0] == 1000.0)            # -1000 * weight (contacts.py:132-133)
    assert np.array_equal(s1[~over, 0], s0[~over, 0])


def test_small_zone_radius(small_system):
    """radius below the pair cutoffs: zone membership gates ligand-receptor pairs too."""
    from gaudi_b200.engine import Engine
    lig, rec, _, _ = small_system
    params = M.ContactParams(radius=2.5)
    rec2 = M.build_receptor(rec.xyz, rec.radius, rec.flags, rec.bonds, params, centre=rec.centre)
    eng = Engine(lig, rec2, params, device=0)
    tors, xf = synth.make_alleles(24, lig.n_allele_torsions, seed=13)
    s, c, x = eng.evaluate(tors, xf, want_xyz=True)
    s = s.cpu().numpy(); x = x.cpu().numpy().astype(np.float64) + eng.centre
    for p in range(24):
        res = OA.contacts(lig, rec2, params, x[p])
        assert score_close(s[p, 0], res.clash, res.clash_slack), (p, s[p], res.clash)
        assert score_close(s[p, 1], res.hydro, res.hydro_slack), (p, s[p], res.hydro)


def test_rmsd_matrix(small_system, torch_cuda):
    lig, rec, params, eng = small_system
    from gaudi_b200.engine import bits_to_bool
    conf = synth.make_conformers(n=333, n_atoms=60, seed=4)
    ref = OA.rmsd_matrix(conf)
    out, bits = eng.rmsd_matrix(conf, threshold=2.5)
    got = out.cpu().numpy()
    assert np.abs(got - ref).max() < RMSD_TOL
    sim = bits_to_bool(bits, 333)
    clear = np.abs(ref - 2.5) > 1e-4
    assert np.array_equal(sim[clear], (ref <= 2.5)[clear])
    # row block == the same rows of the full matrix (row-block sharding)
    out2, bits2 = eng.rmsd_matrix(conf, row0=100, row1=231, threshold=2.5)
    assert torch_cuda.equal(out2, out[100:231]) and torch_cuda.equal(bits2, bits[100:231])
    # bitmask-only calls compute each off-diagonal tile once and mirror it: same words
    _, b_only = eng.rmsd_matrix(conf, threshold=2.5, values=False)
    assert torch_cuda.equal(b_only, bits)
    big = synth.make_conformers(n=1500, n_atoms=60, seed=6)
    _, b_full = eng.rmsd_matrix(big, threshold=2.5)                       # values + bits: no mirroring
    _, b_sym = eng.rmsd_matrix(big, threshold=2.5, values=False)
    assert torch_cuda.equal(b_sym, b_full)
    for r0, r1 in ((128, 1500), (256, 777), (100, 900)):                  # row blocks (multi-GPU shards)
        _, b_blk = eng.rmsd_matrix(big, row0=r0, row1=r1, threshold=2.5, values=False)
        assert torch_cuda.equal(b_blk, b_full[r0:r1]), (r0, r1)
    # partner lists compacted on the device == the set bits of the bitmask, self excluded
    for r0, r1 in ((0, 1500), (100, 900)):
        ptr, idx = eng.similar_lists(b_full[r0:r1], 1500, row0=r0)
        want = bits_to_bool(b_full[r0:r1], 1500)
        want[np.arange(r1 - r0), np.arange(r0, r1)] = False
        assert np.array_equal(ptr, np.concatenate([[0], np.cumsum(want.sum(axis=1))]))
        assert np.array_equal(idx, np.nonzero(want)[1])
    # properties at full size: symmetric, zero diagonal
    big = synth.make_conformers(n=2048, n_atoms=60, seed=5)
    o, _ = eng.rmsd_matrix(big, threshold=2.5, bits=False)
    assert torch_cuda.equal(o, o.T.contiguous())
    assert float(o.diagonal().abs().max()) == 0.0


def test_large_receptor(torch_cuda):
    """120,000 receptor atoms (a solvated complex): per-individual leaf lists and zone words of the
    one-kernel path are then 22 KB of shared memory per pose; warp-per-individual and
    CTA-per-individual variants against the tiled two-kernel path, bit for bit."""
    from gaudi_b200.engine import Engine
    params = M.ContactParams()
    lig = synth.make_ligand(40, 10, seed=1)
    rec = synth.make_receptor(120000, seed=2, params=params)
    fused = Engine(lig, rec, params, device=0)
    tiled = Engine(lig, rec, M.ContactParams(tiled=True), device=0)
    for pop in (300, 3000):
        tors, xf = synth.make_alleles(pop, lig.n_allele_torsions, seed=pop, radius=20.0)
        s1, c1, _ = fused.evaluate(tors, xf)
        s2, c2, _ = tiled.evaluate(tors, xf)
        assert torch_cuda.equal(s1, s2) and torch_cuda.equal(c1, c2)
        assert int((c1.sum(dim=1) > 0).sum()) > pop // 2          # poses do touch the receptor
    fused.close(); tiled.close()


def test_nondominated_fronts_on_device(small_system, torch_cuda):
    """gmm_nondominated_fronts against the definition (peel the non-dominated layer, repeat) for
    2-5 objectives with plenty of ties and duplicates, and against the O(N log N) two-objective
    host sort at a size where peeling is out of reach."""
    from gaudi_b200 import tensor_ga as TG, _lib
    lig, rec, params, eng = small_system
    rng = np.random.default_rng(21)

    def peel(w):
        level = np.full(len(w), -1)
        rem, lv = np.arange(len(w)), 0
        while len(rem):
            sub = w[rem]
            dom = np.zeros(len(sub), bool)
            for s0 in range(0, len(sub), 256):
                blk = sub[s0:s0 + 256]
                dom[s0:s0 + 256] = ((sub[None] >= blk[:, None]).all(2) & (sub[None] > blk[:, None]).any(2)).any(1)
            level[rem[~dom]] = lv
            rem, lv = rem[dom], lv + 1
        return level

    for n, m in ((1, 3), (2, 2), (511, 3), (512, 4), (513, 3), (1700, 3), (1300, 5), (900, 1)):
        w = np.round(rng.normal(size=(n, m)), 1)              # one decimal: ties in every column
        w[n // 2:n // 2 + 5] = w[0]                            # exact duplicates of one row
        assert np.array_equal(eng.nondominated_fronts(w), peel(w)), (n, m)
    w = np.round(np.stack([rng.exponential(50, 70000), -rng.exponential(30, 70000)], 1), 3)
    assert np.array_equal(eng.nondominated_fronts(w), TG.nondominated_fronts(w))
    # rows that are not sorted are refused, not mis-sorted
    bad = torch_cuda.as_tensor(np.array([[0.0, 1.0], [1.0, 0.0], [0.5, 0.5]])).cuda()
    out = torch_cuda.empty(3, dtype=torch_cuda.int32, device="cuda")
    import ctypes as C
    rc = eng._L.gmm_nondominated_fronts(eng._h, 3, 2, C.c_void_p(bad.data_ptr()), C.c_void_p(out.data_ptr()), None)
    assert rc == _lib.GMM_EINVAL


def test_similarity_bitmask_tensor_core_path_edges(small_system, torch_cuda):
    """The tensor-core bitmask kernel decides most pairs from an fp16 product and re-decides the rest
    exactly; the cases here push everything through the re-decision machinery: every pair at the
    threshold (queue overflow), coordinates that do not fit fp16 (nothing pre-decided), identical
    conformers, sizes that are multiples of nothing, few atoms (FP32 kernel)."""
    from gaudi_b200.engine import Engine, bits_to_bool
    lig, rec, params, eng = small_system
    simt = Engine(lig, rec, M.ContactParams(rmsd_simt=True), device=0)
    rng = np.random.default_rng(12)

    def both(conf, thr):
        conf = torch_cuda.as_tensor(np.ascontiguousarray(conf, np.float32)).cuda()
        _, a = eng.rmsd_matrix(conf, threshold=thr, values=False)
        _, b = simt.rmsd_matrix(conf, threshold=thr, values=False)
        assert torch_cuda.equal(a, b)
        return bits_to_bool(a, conf.shape[0])

    base = rng.normal(scale=4.0, size=(1, 47, 3))
    # every pair's RMSD within a few 1e-4 of the threshold: sqrt(2) * sigma * ... = 2.5
    noise = rng.normal(size=(700, 47, 3))
    noise *= 2.5 / np.sqrt(2.0) / np.sqrt((noise ** 2).sum(axis=(1, 2), keepdims=True) / 47)
    sim = both(base + noise, 2.5)
    frac = sim.mean()
    assert 0.02 < frac < 0.98                       # genuinely undecided population, both outcomes occur
    # far from the origin and spread over kilometres: the fp16 copy overflows, all pairs exact
    far = base + rng.normal(scale=3.0e5, size=(300, 1, 3)) + rng.normal(scale=1.0, size=(300, 47, 3))
    far[1] = far[0]
    sim = both(far, 2.5)
    assert sim[0, 1] and sim[1, 0] and sim.sum() < 300 * 300
    # identical conformers: everything similar, including threshold 0 (rmsd == 0 is "not greater")
    same = np.repeat(base, 130, axis=0)
    assert both(same, 2.5).all() and both(same, 0.0).all()
    # ragged: N = 1, N = 129 (one full tile + one row), A = 9 (K = 27), A = 3 (FP32 kernel)
    for n, a_ in ((1, 20), (129, 9), (257, 33), (40, 3)):
        conf = rng.normal(scale=2.0, size=(n, a_, 3))
        sim = both(conf, 2.5)
        ref = OA.rmsd_matrix(conf.astype(np.float32).astype(np.float64))
        clear = np.abs(ref - 2.5) > 1e-4
        assert np.array_equal(sim[clear], (ref <= 2.5)[clear])


@pytest.mark.parametrize("n_lig, n_tors, n_rec, pop", [
    (1, 0, 7, 3),          # a single mobile atom, receptor smaller than one leaf
    (5, 0, 33, 5),         # just over one leaf
    (40, 10, 129, 2665),   # just over one 128-atom tile; population just over the warp/CTA switch
    (70, 12, 1000, 9),     # mobile set larger than two lane-rounds
])
def test_ragged_sizes(torch_cuda, n_lig, n_tors, n_rec, pop):
    """Sizes that do not line up with any tile, leaf, warp or team: fused == tiled bit for bit,
    and both equal the oracle."""
    from gaudi_b200.engine import Engine
    params = M.ContactParams()
    if n_lig >= 12:
        lig = synth.make_ligand(n_lig, n_tors, seed=3)
    else:
        xyz0 = np.array([[0.0, 0.0, 0.0], [1.5, 0.0, 0.0], [2.2, 1.3, 0.0], [3.7, 1.3, 0.2], [4.4, 2.6, 0.1]])[:n_lig]
        bonds = [(i, i + 1) for i in range(n_lig - 1)]
        flags = np.full(n_lig, M.F_HYDROPHOBIC | M.F_SEED | M.F_XFORM, np.uint8)
        lig = M.build_mobile(xyz0, np.full(n_lig, 1.7), flags, bonds, [], origin_atom=0)
    rec = synth.make_receptor(n_rec, seed=4, cavity=2.0, params=params)
    tors, xf = synth.make_alleles(pop, lig.n_allele_torsions, seed=5, radius=3.0)
    eng = Engine(lig, rec, params, device=0)
    s1, c1, x1 = eng.evaluate(tors, xf, want_xyz=True)
    tiled = Engine(lig, rec, M.ContactParams(tiled=True), device=0)
    s2, c2 = tiled.contacts(x1)
    assert torch_cuda.equal(s1, s2) and torch_cuda.equal(c1, c2)
    s1 = s1.cpu().numpy(); x = x1.cpu().numpy().astype(np.float64) + eng.centre
    for p in range(0, pop, max(1, pop // 6)):
        res = OA.contacts(lig, rec, params, x[p])
        assert score_close(s1[p, 0], res.clash, res.clash_slack), (p, s1[p], res.clash)
        assert score_close(s1[p, 1], res.hydro, res.hydro_slack), (p, s1[p], res.hydro)
    eng.close(); tiled.close()


def test_zone_radius_zero_and_huge(small_system, torch_cuda):
    """radius = 0: the zone is the probe itself (contacts.py:199-213 with an empty
    neighbourhood); radius = 30: the zone is everything."""
    from gaudi_b200.engine import Engine
    lig, rec, _, _ = small_system
    tors, xf = synth.make_alleles(12, lig.n_allele_torsions, seed=23)
    for radius in (0.0, 30.0):
        params = M.ContactParams(radius=radius)
        rec2 = M.build_receptor(rec.xyz[:1500], rec.radius[:1500], rec.flags[:1500],
                                rec.bonds[(rec.bonds < 1500).all(axis=1)], params, centre=rec.centre)
        eng = Engine(lig, rec2, params, device=0)
        s, c, x = eng.evaluate(tors, xf, want_xyz=True)
        s = s.cpu().numpy(); x = x.cpu().numpy().astype(np.float64) + eng.centre
        for p in range(0, 12, 3):
            res = OA.contacts(lig, rec2, params, x[p])
            assert score_close(s[p, 0], res.clash, res.clash_slack), (radius, p, s[p], res.clash)
            assert score_close(s[p, 1], res.hydro, res.hydro_slack), (radius, p, s[p], res.hydro)
        eng.close()


def test_full_size_baseline_config(torch_cuda):
    """BASELINE.json configs[2] at full size (65,536 poses x 20,000-atom receptor): culled ==
    dense bit for bit over the whole population, a sample of it equals the oracle, and the
    population evaluated in 8 shards (what 8 GPUs do) gives the same rows."""
    from gaudi_b200.engine import Engine
    params = M.ContactParams()
    lig = synth.make_ligand(40, 10, seed=1)
    rec = synth.make_receptor(20000, seed=2, params=params)
    tors, xf = synth.make_alleles(65536, 10, seed=3)
    eng = Engine(lig, rec, params, device=0)
    dense = Engine(lig, rec, M.ContactParams(cull=False), device=0)
    eng.set_profiling(True); dense.set_profiling(True)
    s1, c1, x1 = eng.evaluate(tors, xf, want_xyz=True)
    s2, c2 = dense.contacts(x1)
    assert torch_cuda.equal(s1, s2) and torch_cuda.equal(c1, c2)
    assert 0 < eng.last_pair_tests() * 10 < dense.last_pair_tests()
    eng.set_profiling(False)
    parts = [eng.evaluate(tors[k * 8192:(k + 1) * 8192], xf[k * 8192:(k + 1) * 8192])[0] for k in range(8)]
    assert torch_cuda.equal(torch_cuda.cat(parts), s1)
    s = s1.cpu().numpy(); x = x1.cpu().numpy().astype(np.float64) + eng.centre
    for p in range(0, 65536, 2731):                      # 24 individuals spread over the population
        assert np.abs(OA.express(lig, tors[p], xf[p]) - x[p]).max() < XYZ_TOL
        res = OA.contacts(lig, rec, params, x[p])
        assert score_close(s[p, 0], res.clash, res.clash_slack), (p, s[p], res.clash)
        assert score_close(s[p, 1], res.hydro, res.hydro_slack), (p, s[p], res.hydro)
    assert np.isfinite(s).all()


def test_api_errors_and_reserve(small_system, torch_cuda):
    import ctypes as C
    from gaudi_b200 import _lib
    lig, rec, params, eng = small_system
    L = _lib.lib()
    eng.reserve(10000)
    tors, xf = synth.make_alleles(16, lig.n_allele_torsions, seed=1)
    # a model with torsions needs the torsion alleles
    with pytest.raises(_lib.GmmError) as e:
        eng.express(None, xf)
    assert e.value.code == _lib.GMM_EINVAL
    # parameters cannot change once the receptor's static pair values exist
    assert L.gmm_set_params(eng._h, 5.0, 0.6, -0.4, 0.2, 2.1, 0.0, 1.0, 0) == _lib.GMM_ESTATE
    # a fresh context refuses to score before it has a model
    h = C.c_void_p()
    assert L.gmm_create(C.byref(h), 0) == 0
    out = torch_cuda.zeros((1, 2), dtype=torch_cuda.float64, device="cuda")
    x = torch_cuda.zeros((1, 3), dtype=torch_cuda.float32, device="cuda")
    assert L.gmm_contacts(h, 1, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), None, None) == _lib.GMM_ESTATE
    assert L.gmm_create(C.byref(C.c_void_p()), 9999) == _lib.GMM_EINVAL
    assert L.gmm_destroy(h) == 0


def test_evaluate_host_pinned_chunked(small_system, torch_cuda):
    """Page-locked host buffers take the chunked upload/compute-overlap path; same bits."""
    lig, rec, params, eng = small_system
    pop = 20001                                            # ragged: chunks of 2500 and 17501
    tors, xf = synth.make_alleles(pop, lig.n_allele_torsions, seed=29)
    want_s, want_c, _ = eng.evaluate(tors, xf)
    t_pin = torch_cuda.as_tensor(tors).pin_memory()
    x_pin = torch_cuda.as_tensor(xf).pin_memory()
    s_pin = torch_cuda.full((pop, 2), float("nan"), dtype=torch_cuda.float64).pin_memory()
    c_pin = torch_cuda.full((pop, 2), -1, dtype=torch_cuda.int32).pin_memory()
    for _ in range(2):
        eng.evaluate_host(t_pin, x_pin, scores_out=s_pin, counts_out=c_pin)
        assert torch_cuda.equal(s_pin, want_s.cpu()) and torch_cuda.equal(c_pin, want_c.cpu())
        s_pin.fill_(float("nan"))


def test_whole_population_against_c_oracle(torch_cuda):
    """EVERY individual of BASELINE.json configs[1] (4,096 poses, 5,000-atom receptor) and a 16,384
    slice of configs[2] (20,000-atom receptor) against the oracle's C restatement (itself tied to
    the numpy oracle and the reference's goldens by tests/test_c_oracle.py).  Individuals whose
    score differs by more than 1e-4 relative must be explained by pairs within 1e-4 A of a cutoff
    (north_star's exemption), which the numpy oracle adapter quantifies."""
    import os, sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
    import c_oracle
    from gaudi_b200.engine import Engine
    for n_rec, pop in ((5000, 4096), (20000, 16384)):
        params = M.ContactParams()
        lig = synth.make_ligand(40, 10, seed=1)
        rec = synth.make_receptor(n_rec, seed=2, params=params)
        tors, xf = synth.make_alleles(pop, 10, seed=3)
        eng = Engine(lig, rec, params, device=0)
        s, c, x = eng.evaluate(tors, xf, want_xyz=True)
        s = s.cpu().numpy(); c = c.cpu().numpy(); x = x.cpu().numpy().astype(np.float64) + eng.centre
        ref_s, ref_c, ref_x = c_oracle.COracle(lig, rec, params).evaluate(tors, xf, want_xyz=True)
        assert np.abs(x - ref_x).max() < XYZ_TOL
        rel = np.abs(s - ref_s) / np.maximum(1.0, np.abs(ref_s))
        bad = np.flatnonzero((rel > SCORE_RTOL).any(axis=1) | (c != ref_c).any(axis=1))
        assert len(bad) < 0.02 * pop, len(bad)            # borderline poses are rare
        for p in bad[:40]:
            res = OA.contacts(lig, rec, params, ref_x[p])
            assert res.borderline, (p, s[p], ref_s[p])
            assert score_close(s[p, 0], res.clash, res.clash_slack), (p, s[p], res.clash, res.clash_slack)
            assert score_close(s[p, 1], res.hydro, res.hydro_slack), (p, s[p], res.hydro, res.hydro_slack)
        eng.close()
