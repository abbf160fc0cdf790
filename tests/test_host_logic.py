"""Host-side logic that needs no GPU: NSGA-II bookkeeping, the greedy niching pass against the
oracle's restatement, allele plumbing of the gene plugins, sharding arithmetic, and the N > 1
gather path over gloo (world size 2) with the CPU oracle standing in for the per-shard compute."""
import os
import random
import sys

import numpy as np
import pytest

import oracle_adapter as OA
from gaudi_b200 import deaplite, parallel
from gaudi_b200.algorithms import BatchedMap, niching

O = OA.oracle()


class _Ind(object):
    def __init__(self, values, weights=(-1.0, -1.0)):
        self.fitness = deaplite.Fitness(weights)
        self.fitness.values = values


def test_fitness_and_nsga2():
    a, b, c = _Ind((1.0, 1.0)), _Ind((2.0, 2.0)), _Ind((0.5, 3.0))
    assert a.fitness.dominates(b.fitness) and not b.fitness.dominates(a.fitness)
    assert not a.fitness.dominates(c.fitness) and not c.fitness.dominates(a.fitness)
    chosen = deaplite.sel_nsga2([a, b, c], 2)
    assert set(map(id, chosen)) == {id(a), id(c)}
    fronts = deaplite.sort_nondominated([a, b, c], 3)
    assert [len(f) for f in fronts] == [2, 1]
    pf = deaplite.ParetoFront(lambda x, y: False)
    pf.update([a, b, c])
    assert len(pf) == 2


def test_niching_matches_oracle_restatement():
    rng = random.Random(5)
    inds = [_Ind((rng.random(), rng.random())) for _ in range(40)]
    sim = {(i, j) for i in range(40) for j in range(40) if i != j and rng.random() < 0.15}
    sim |= {(j, i) for i, j in sim}
    index = {id(x): k for k, x in enumerate(inds)}
    kept = niching(inds, lambda a, b: (index[id(a)], index[id(b)]) in sim)
    want = O.greedy_niching(list(range(40)), lambda i, j: (i, j) in sim,
                            lambda i, j: inds[i].fitness <= inds[j].fitness)
    assert [index[id(x)] for x in kept] == want


def test_variation_operators_respect_bounds():
    random.seed(3)
    a, b = [10.0, -170.0, 33.3], [-20.0, 160.0, 90.0]
    deaplite.cx_simulated_binary_bounded(a, b, 5.0, -180.0, 180.0)
    assert all(-180.0 <= v <= 180.0 for v in a + b)
    m, = deaplite.mut_polynomial_bounded([0.0, 179.0, -179.0], 5.0, -180.0, 180.0, 1.0)
    assert all(-180.0 <= v <= 180.0 for v in m)
    x, y = deaplite.cx_two_point([1, 2, 3, 4], [5, 6, 7, 8])
    assert sorted(x + y) == [1, 2, 3, 4, 5, 6, 7, 8]


def test_batched_map_dispatch():
    class Env(object):
        def evaluate_population(self, inds):
            return [[len(inds)]] * len(inds)

    def evaluate(ind):
        raise AssertionError("serial path must not run")
    bm = BatchedMap(Env(), evaluate)
    assert bm(evaluate, [1, 2, 3]) == [[3], [3], [3]]
    assert bm(abs, [-1, 2]) == [1, 2]


def test_search_allele_product():
    from gaudi_b200.genes import search as S
    random.seed(7)
    t = S.random_translation((1.0, 2.0, 3.0), 4.0)
    r = S.random_rotation()
    prod = S.multiply_matrices(t, r)
    assert np.allclose(prod[:, :3] @ prod[:, :3].T, np.eye(3), atol=1e-12)
    assert np.allclose(prod[:, 3], [row[3] for row in t])
    assert np.linalg.norm(prod[:, 3] - np.array([1.0, 2.0, 3.0])) <= 4.0
    assert np.allclose(prod, O.multiply_3x4(t, r))
    mid = S.interpolate_xforms(prod, (0, 0, 0), prod, 0.5)
    assert np.allclose(mid, prod, atol=1e-12)


def test_shard_bounds():
    for n in (0, 1, 7, 65536, 65537):
        for w in (1, 2, 3, 8):
            b = parallel.shard_bounds(n, w)
            assert b[0] == 0 and b[-1] == n and all(b[i] <= b[i + 1] for i in range(w))
            assert max(b[i + 1] - b[i] for i in range(w)) - min(b[i + 1] - b[i] for i in range(w)) <= 1


def _gloo_worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gaudi_b200 import model as M, synth
    params = M.ContactParams()
    lig = synth.make_ligand(12, 3, seed=1)
    rec = synth.make_receptor(300, seed=2, params=params)
    tors, xf = synth.make_alleles(9, 3, seed=3)          # 9 individuals over 2 ranks: ragged

    def shard(t, x):          # the CPU oracle stands in for the GPU on this rank's slice
        res = []
        for k in range(t.shape[0]):
            r = OA.contacts(lig, rec, params, OA.express(lig, t[k], x[k]), with_slack=False)
            res.append((r.clash, r.hydro))
        return np.array(res).reshape(-1, 2)
    full = parallel.sharded_evaluate(shard, (tors, xf))
    conf = np.stack([OA.express(lig, tors[k], xf[k]) for k in range(9)]).astype(np.float32)
    b = parallel.shard_bounds(9, world)

    def rows(allc, r0, r1):
        ref = OA.rmsd_matrix(allc.numpy().astype(np.float64), rows=range(r0, r1)) <= 2.5
        words = np.packbits(np.pad(ref, ((0, 0), (0, 32 - ref.shape[1] % 32)))[:, :32 * ((ref.shape[1] + 31) // 32)],
                            axis=1, bitorder="little").view(np.uint32).astype(np.int64)
        return torch.as_tensor(words.astype(np.uint32).view(np.int32))
    bits = parallel.sharded_rmsd_bits(rows, torch.as_tensor(conf[b[rank]:b[rank + 1]]))
    # block sizes known to the caller (no size exchange), and equal blocks (gathered in place)
    known = parallel.sharded_rmsd_bits(rows, torch.as_tensor(conf[b[rank]:b[rank + 1]]),
                                       counts=[b[r + 1] - b[r] for r in range(world)])
    assert torch.equal(known, bits)
    b8 = parallel.shard_bounds(8, world)
    even = parallel.sharded_rmsd_bits(rows, torch.as_tensor(conf[:8][b8[rank]:b8[rank + 1]]),
                                      counts=[b8[r + 1] - b8[r] for r in range(world)])
    assert torch.equal(even, rows(torch.as_tensor(conf[:8]), 0, 8))
    if rank == 0:
        want = shard(tors, xf)
        ref = OA.rmsd_matrix(conf.astype(np.float64)) <= 2.5
        from gaudi_b200.engine import bits_to_bool
        np.save(out, np.array([float(np.abs(full.numpy() - want).max()),
                               float((bits_to_bool(bits.numpy(), 9) != ref).sum())]))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_gather_world2_gloo(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "res.npy")
    port = 29500 + random.randint(0, 2000)
    mp.spawn(_gloo_worker, args=(2, port, out), nprocs=2, join=True)
    res = np.load(out)
    assert res[0] == 0.0          # gathered fitness == single-process fitness, bit for bit
    assert res[1] == 0.0          # row-block sharded similarity bitmask == full matrix


def _host_sharded_worker(rank, world, port, out):
    import torch.distributed as dist
    from collections import OrderedDict
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    spec = OrderedDict([("x", ((3,), np.float64)), ("y", ((2,), np.float64)), ("who", ((), np.int32))])
    table = parallel.HostSharded(spec, 50, timeout=30.0)
    a = table.arrays

    def block(lo, hi):          # stands in for this rank's GPU: a function of its rows only
        a["y"][lo:hi, 0] = a["x"][lo:hi].sum(axis=1)
        a["y"][lo:hi, 1] = a["x"][lo:hi, 0] * 2.0
        a["who"][lo:hi] = rank

    if rank == 0:
        rng = np.random.default_rng(3)
        worst, owners = 0.0, []
        for n in (50, 7, 1, 0, 33):                      # full table, ragged, fewer rows than ranks, empty
            a["x"][:n] = rng.normal(size=(n, 3))
            a["y"][:] = np.nan
            b = table.run(n, block)
            assert b == parallel.shard_bounds(n, world)
            want = np.stack([a["x"][:n].sum(axis=1), a["x"][:n, 0] * 2.0], axis=1) if n else np.zeros((0, 2))
            worst = max(worst, float(np.abs(a["y"][:n] - want).max()) if n else 0.0)
            owners.append([int(np.sum(a["who"][:n] == r)) for r in range(world)])
        table.shutdown()
        np.save(out, np.array([worst] + [float(v) for row in owners for v in row]))
    else:
        table.serve(block)                               # returns when rank 0 shuts the job down
    table.close()
    dist.barrier()
    dist.destroy_process_group()


def test_host_sharded_table_world2_gloo(tmp_path):
    """The shared population table behind multi-GPU ``toolbox.map``: rank 0 drives steps, the other
    rank serves until shutdown; every row is computed by the rank that owns it."""
    import torch.multiprocessing as mp
    out = str(tmp_path / "hs.npy")
    port = 31500 + random.randint(0, 2000)
    mp.spawn(_host_sharded_worker, args=(2, port, out), nprocs=2, join=True)
    res = np.load(out)
    assert res[0] == 0.0
    assert res[1:].reshape(-1, 2).tolist() == [[25, 25], [3, 4], [0, 1], [0, 0], [16, 17]]


# ---- tensorised GA bookkeeping (gaudi_b200/tensor_ga.py) against the per-object versions ----------
def test_tensor_nsga2_matches_deaplite():
    from gaudi_b200 import tensor_ga as TG
    rng = np.random.default_rng(4)
    for nobj in (2, 3):
        vals = rng.random((60, nobj))
        vals[7] = vals[3]                              # duplicated fitness shares a front
        weights = (-1.0,) * nobj
        inds = [_Ind(tuple(v), weights) for v in vals]
        index = {id(x): i for i, x in enumerate(inds)}
        level = TG.nondominated_fronts(vals * np.array(weights))
        for k in (60, 25, 1):
            want = [index[id(x)] for x in deaplite.sel_nsga2(inds, k)]
            got = TG.sel_nsga2(vals, weights, k).tolist()
            assert len(got) == len(set(got)) == k

            def key(sel):      # ties in crowding distance are broken arbitrarily by either version
                out = []
                for i in sel:
                    members = np.flatnonzero(level == level[i])
                    cd = TG.crowding_distance(vals[members])[list(members).index(i)]
                    out.append((int(level[i]), float(np.round(cd, 9))))
                return sorted(out)
            assert key(got) == key(want)
        fronts = deaplite.sort_nondominated(inds, 60)
        for lv, front in enumerate(fronts):
            assert all(level[index[id(x)]] == lv for x in front)


def test_tensor_niching_matches_reference_loop():
    from gaudi_b200 import tensor_ga as TG
    rng = np.random.default_rng(9)
    n = 70
    vals = np.round(rng.random((n, 2)), 1)             # coarse values: plenty of fitness ties
    sim = rng.random((n, n)) < 0.12
    sim = sim | sim.T
    np.fill_diagonal(sim, True)
    inds = [_Ind(tuple(v)) for v in vals]
    index = {id(x): k for k, x in enumerate(inds)}
    want = [index[id(x)] for x in niching(inds, lambda a, b: bool(sim[index[id(a)], index[id(b)]]))]
    w = vals * np.array([-1.0, -1.0])
    got = TG.niching(sim, w).tolist()
    assert got == want
    # same pass on the packed words gmm_rmsd_matrix writes, and on partner lists (gmm_niching_host)
    padded = np.zeros((n, 96), np.uint8)
    padded[:, :n] = sim
    words = np.packbits(padded, axis=1, bitorder="little").view(np.uint32).view(np.int32)
    assert TG.niching(words, w).tolist() == want
    off = sim.copy()
    np.fill_diagonal(off, False)
    ptr = np.concatenate([[0], np.cumsum(off.sum(axis=1))])
    assert TG.niching((ptr, np.nonzero(off)[1]), w).tolist() == want
    assert TG.niching((np.zeros(1, np.int64), np.zeros(0, np.int32)), np.zeros((0, 2))).tolist() == []


def test_two_objective_fronts_match_peeling():
    from gaudi_b200 import tensor_ga as TG
    rng = np.random.default_rng(4)
    for n in (1, 2, 37, 400):
        w = np.round(rng.normal(size=(n, 2)), 1)
        uniq, inv = np.unique(w, axis=0, return_inverse=True)
        level = np.full(len(uniq), -1)
        rem, lv = np.arange(len(uniq)), 0
        while len(rem):                                   # definition: peel non-dominated layers
            sub = uniq[rem]
            dom = ((sub[None] >= sub[:, None]).all(2) & (sub[None] > sub[:, None]).any(2)).any(1)
            level[rem[~dom]] = lv
            rem, lv = rem[dom], lv + 1
        assert np.array_equal(TG.nondominated_fronts(w), level[inv.reshape(-1)])


def test_tensor_variation_properties():
    from gaudi_b200 import tensor_ga as TG
    spec = dict(n_torsions=6, flexibility=300.0, precision=1, cx_eta=5.0, mut_eta=5.0, mut_indpb=0.75,
                search_center=(1.0, 2.0, 3.0), search_radius=4.0, rotate=True, interpolation=0.5, n_rot_res=3)
    ga = TG.TensorGA(spec, (-1.0, -1.0), None, None, seed=1)
    pop = ga.random_population(200)
    pop["fitness"][:] = 1.0
    pop["coords"] = np.zeros((200, 5, 3), np.float32)
    kids = ga.var_or(pop, 500, 0.5, 0.4)
    assert kids["torsions"].shape == (500, 6) and np.abs(kids["torsions"]).max() <= 150.0
    assert np.allclose(kids["torsions"], np.round(kids["torsions"], 1))
    assert np.allclose(np.einsum("nij,nkj->nik", kids["rot"], kids["rot"]), np.eye(3), atol=1e-9)
    assert (np.linalg.norm(kids["trans"] - np.array([1.0, 2.0, 3.0]), axis=1) <= 4.0 + 1e-9).all()
    assert ((kids["rotamers"] >= 0) & (kids["rotamers"] < 1)).all()
    changed = np.isnan(kids["fitness"][:, 0])
    assert 0.8 < changed.mean() < 0.97               # cxpb + mutpb = 0.9 of the children are new
    r = TG.random_rotations(50, np.random.default_rng(0))
    assert np.allclose(TG.slerp(r, r, 0.3), r, atol=1e-9)
    assert np.allclose(TG.quat_to_matrix(TG.matrix_to_quat(r)), r, atol=1e-9)


def test_reference_job_file_is_redirected(tmp_path):
    """A job file written for the reference (``module: gaudi.genes.torsion`` ...) resolves to the
    B200 mirrors; plugins without a mirror are refused by name."""
    from gaudi_b200 import parse
    job = tmp_path / "job.yaml"
    job.write_text("""
ga: {population: 12, generations: 1}
similarity: {module: gaudi.similarity.rmsd, args: [[Ligand], 2.5], kwargs: {}}
genes:
  - {name: Protein, module: gaudi.genes.molecule, path: p.mol2}
  - {name: Ligand, module: gaudi.genes.molecule, path: sub/l.mol2}
  - {name: Torsion, module: gaudi.genes.torsion, target: Ligand, flexibility: 360}
  - {name: Search, module: gaudi.genes.search, target: Ligand/1, radius: 3.0, rotate: true}
objectives:
  - {name: Clashes, module: gaudi.objectives.contacts, which: clashes, probes: [Ligand], radius: 5.0, weight: -1.0}
  - {name: Near, module: gaudi.objectives.distance, probes: [Ligand/1], target: Protein/2, threshold: 3.0, weight: -1.0}
""")
    cfg = parse.Settings(str(job))
    assert [g.module for g in cfg.genes] == ["gaudi_b200.genes.molecule"] * 2 + ["gaudi_b200.genes.torsion",
                                                                                  "gaudi_b200.genes.search"]
    assert [o.module for o in cfg.objectives] == ["gaudi_b200.objectives.contacts", "gaudi_b200.objectives.distance"]
    assert cfg.similarity.module == "gaudi_b200.similarity.rmsd"
    assert cfg.genes[1].path == str(tmp_path / "sub" / "l.mol2")          # relative to the job file
    assert cfg.ga.mu == 1 and cfg.ga.cx_eta == 5                           # reference defaults (parse.py:254-284)
    bad = tmp_path / "bad.yaml"
    bad.write_text("genes:\n  - {name: M, module: gaudi.genes.mutamers, residues: [Protein/1]}\n")
    with pytest.raises(NotImplementedError, match="gaudi.genes.mutamers"):
        parse.Settings(str(bad))


def test_results_writer_round_trip(tmp_path):
    """Individual.write / dump_population (gaudi/base.py:231-266, gaudi/algorithms.py:205-223): one
    zip per individual holding a mol2 per Molecule gene with the EXPRESSED coordinates, the other
    genes' alleles as text and a yaml with the scores, plus the .gaudi-output index; read back, the
    files give the alleles, the scores and the coordinates that went in."""
    import ast
    import zipfile
    import yaml
    import golden_util as G
    from gaudi_b200 import algorithms, base, molio, parse
    random.seed(4)
    cfg = parse.Settings({
        "output": {"path": str(tmp_path / "results"), "name": "job", "verbose": False, "check_every": 0},
        "genes": [{"name": "Protein", "module": "gaudi_b200.genes.molecule", "path": "p.mol2", "mol": G.load("5er1_protein_mol2")},
                  {"name": "Ligand", "module": "gaudi_b200.genes.molecule", "path": "l.mol2", "mol": G.load("5er1_ligand_mol2")},
                  {"name": "Torsion", "module": "gaudi_b200.genes.torsion", "target": "Ligand", "flexibility": 360},
                  {"name": "Search", "module": "gaudi_b200.genes.search", "target": "Ligand/1", "radius": 3.0}],
        "objectives": [{"name": "Clashes", "module": "gaudi_b200.objectives.contacts", "which": "clashes",
                        "probes": ["Ligand"], "radius": 5.0, "weight": -1.0},
                       {"name": "Contacts", "module": "gaudi_b200.objectives.contacts", "which": "hydrophobic",
                        "probes": ["Ligand"], "radius": 5.0, "weight": -1.0}]})
    os.makedirs(cfg.output.path)
    rng = np.random.default_rng(2)
    pop = []
    for k in range(3):
        ind = base.MolecularIndividual(cfg)
        ind.fitness.values = (round(10.0 * k + 0.125, 3), round(-3.5 - k, 3))
        for gene in ind._molecules.values():        # what the batched evaluate leaves behind (post_express)
            gene._expressed_coordinates = gene.mol.xyz + rng.normal(size=gene.mol.xyz.shape)
        pop.append(ind)
    index = algorithms.dump_population(pop, cfg, subdir="check7")
    listing = yaml.safe_load(open(index))
    assert listing["GAUDI.objectives"] == ["Clashes (gaudi_b200.objectives.contacts)", "Contacts (gaudi_b200.objectives.contacts)"]
    assert sorted(listing["GAUDI.results"]) == ["job_000.zip", "job_001.zip", "job_002.zip"]
    for k, ind in enumerate(pop):
        assert listing["GAUDI.results"]["job_%03d.zip" % k] == list(ind.fitness.values)
        with zipfile.ZipFile(os.path.join(cfg.output.path, "check7", "job_%03d.zip" % k)) as z:
            names = sorted(z.namelist())
            assert names == sorted(["Protein_%03d_Protein.mol2" % k, "Ligand_%03d_Ligand.mol2" % k,
                                    "Torsion_%03d_Torsion.txt" % k, "Search_%03d_Search.txt" % k, "Search_%03d.gaudi" % k])
            meta = yaml.safe_load(z.read("Search_%03d.gaudi" % k))
            assert meta["score"] == list(ind.fitness.values) and meta["Ligand"] == "Ligand_%03d_Ligand.mol2" % k
            assert ast.literal_eval(z.read("Torsion_%03d_Torsion.txt" % k).decode()) == list(ind.genes["Torsion"].allele)
            got = ast.literal_eval(z.read("Search_%03d_Search.txt" % k).decode())
            assert np.allclose(np.array(got[0]), np.array(ind.genes["Search"].allele[0]))
            z.extract("Ligand_%03d_Ligand.mol2" % k, str(tmp_path))
        back = molio.read_mol2(str(tmp_path / ("Ligand_%03d_Ligand.mol2" % k)))
        lig = ind.find_molecule("Ligand")
        assert np.abs(back.xyz - lig._expressed_coordinates).max() < 1e-6
        assert np.array_equal(back.bonds, lig.mol.bonds) and back.sybyl == lig.mol.sybyl
    assert not [f for f in os.listdir(os.path.join(cfg.output.path, "check7")) if f.endswith((".mol2", ".txt"))]


def test_coordination_plugin_site_tables():
    """The Coordination plugin's host prep (no GPU): key precedence atom_types > atom_elements > atom_names
    (gaudi/objectives/coordination.py:244-252), residue filters ('*' or a position), the reference's
    defaults and its constructor checks."""
    from types import SimpleNamespace
    import golden_util as G
    from gaudi_b200 import molio
    from gaudi_b200.objectives import coordination as CO
    mol = G.load("4c3w_mol2")
    adj = mol.adjacency()
    nbs = [[] for _ in range(mol.n_atoms)]
    for a, b in mol.bonds:
        nbs[int(a)].append(int(b)); nbs[int(b)].append(int(a))
    system = SimpleNamespace(
        xyz=mol.xyz.astype(np.float64), element=list(mol.element), name=list(mol.name),
        idatm=[molio.idatm_type(mol, i, adj) for i in range(mol.n_atoms)], molecule=["Molecule"] * mol.n_atoms,
        resseq=[int(r) for r in mol.resseq], neighbours=nbs, mobile_row=-np.ones(mol.n_atoms, np.int32),
        locate=lambda spec: int(np.flatnonzero(mol.serial == int(spec[1]))[0]))
    o = CO.enable(probe="Molecule/1005", residues=["Molecule/*"], atom_elements=["C", "N", "O", "S"],
                  geometry="octahedron", weight=-1.0, name="site")
    s = o.site(system)
    assert s["metal"] == int(np.flatnonzero(mol.serial == 1005)[0]) and s["template"].shape == (7, 3)
    assert (s["radius"], s["distance"], s["min_atoms"], s["prevent_intruders"]) == (3.0, 0.0, 1, True)
    want = np.array([e.upper() in ("C", "N", "O", "S") for e in mol.element])
    assert np.array_equal(s["candidate"].astype(bool), want) and not s["candidate"][s["metal"]]
    assert s["geometry"][int(np.flatnonzero(mol.serial == 408)[0])] == 3          # carboxylate O: planar
    assert abs(s["bond_radius"][s["metal"]] - 1.33) < 1e-12                       # vanadium
    # one residue only; atom_types wins over atom_elements, which wins over atom_names
    o = CO.enable(probe=["Molecule", 1005], residues=[["Molecule", 52]], atom_types=["O2-"], atom_elements=["N"],
                  atom_names=["CA"], weight=-1.0, name="site")
    s = o.site(system)
    picked = np.flatnonzero(s["candidate"])
    assert len(picked) and all(int(mol.resseq[i]) == 52 and system.idatm[i] == "O2-" for i in picked)
    o = CO.enable(probe="Molecule/1005", residues=["Molecule/130"], atom_names=["O2", "N2"], weight=-1.0, name="site")
    picked = np.flatnonzero(o.site(system)["candidate"])
    assert sorted(mol.name[i] for i in picked) == ["N2", "O2"]
    with pytest.raises(ValueError):
        CO.enable(probe="Molecule/1005", residues=["Molecule/*"], weight=-1.0, name="site")      # no atom filter
    with pytest.raises(ValueError):
        CO.enable(probe="Molecule/1005", residues=["Molecule/*"], atom_elements=["O"], bogus=1, name="site")
    assert sorted(CO.GEOMETRIES) == sorted(["cube", "hexagonal bipyramid", "linear", "octahedron", "pentagonal bipyramid",
                                            "square planar", "square pyramid", "tetrahedral", "trigonal planar",
                                            "trigonal bipyramid", "trigonal prism"])


def test_molecule_write_combined(tmp_path):
    """Molecule.write(combined_with=[...]) (gaudi/genes/molecule.py:252-258): every molecule's atoms in
    one file, renumbered, bonds carried over; mol2 and pdb."""
    import golden_util as G
    from gaudi_b200 import molio
    lig, prot = G.load("5er1_ligand_mol2"), G.load("5er1_protein_mol2")
    moved = lig.xyz + np.array([1.0, -2.0, 0.5])
    both = molio.combine([(lig, moved), (prot, None)])
    assert both.n_atoms == lig.n_atoms + prot.n_atoms and len(both.bonds) == len(lig.bonds) + len(prot.bonds)
    assert list(both.serial) == list(range(1, both.n_atoms + 1))
    path = molio.write_mol2(both, str(tmp_path / "both.mol2"))
    back = molio.read_mol2(path)
    assert back.n_atoms == both.n_atoms and len(back.bonds) == len(both.bonds)
    assert np.abs(back.xyz[:lig.n_atoms] - moved).max() < 1e-5
    assert np.abs(back.xyz[lig.n_atoms:] - prot.xyz).max() < 1e-5
    # the first protein bond joins the same two protein atoms as before, shifted by the ligand's atoms
    assert tuple(int(v) for v in back.bonds[len(lig.bonds)]) == tuple(int(v) + lig.n_atoms for v in prot.bonds[0])
    pdb = molio.read_pdb(molio.write_pdb(both, str(tmp_path / "both.pdb")))
    assert pdb.n_atoms == both.n_atoms and np.abs(pdb.xyz[:lig.n_atoms] - moved).max() < 2e-3
