"""The reference-facing plugin path on the GPU: gene / objective plugins loaded by dotted module
path, batched ``toolbox.map``, the (mu + lambda) loop with GPU-computed similarity, flexible
side chains (Rotamers) — each checked against the CPU oracle.  BASELINE.json configs[0]
(tests/data ligand + receptor, Torsion + Search, Contacts, pop 100 x 10 generations) runs here
from the committed fixtures."""
import random

import numpy as np
import pytest

import golden_util as G
import oracle_adapter as OA
from gaudi_b200 import parse

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cuda(built_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch


def job(population=100, generations=10, rotamers=None):
    genes = [
        {"name": "Protein", "module": "gaudi_b200.genes.molecule", "path": "5er1_protein.mol2",
         "mol": G.load("5er1_protein_mol2")},
        {"name": "Ligand", "module": "gaudi_b200.genes.molecule", "path": "5er1_ligand.mol2",
         "mol": G.load("5er1_ligand_mol2")},
    ]
    if rotamers:
        genes.append({"name": "Rotamers", "module": "gaudi_b200.genes.rotamers", "residues": rotamers[0],
                      "library": rotamers[1], "with_original": True})
    genes += [
        {"name": "Torsion", "module": "gaudi_b200.genes.torsion", "target": "Ligand", "flexibility": 360},
        {"name": "Search", "module": "gaudi_b200.genes.search", "target": "Ligand/1", "radius": 3.0, "rotate": True},
    ]
    return parse.Settings({
        "output": {"verbose": False, "precision": 3, "check_every": 0},
        "ga": {"population": population, "generations": generations, "mu": 1.0, "lambda_": 1.0},
        "genes": genes,
        "objectives": [
            {"name": "Clashes", "module": "gaudi_b200.objectives.contacts", "which": "clashes",
             "probes": ["Ligand"], "radius": 5.0, "weight": -1.0},
            {"name": "Contacts", "module": "gaudi_b200.objectives.contacts", "which": "hydrophobic",
             "probes": ["Ligand"], "radius": 5.0, "weight": -1.0}],
        "similarity": {"module": "gaudi_b200.similarity.rmsd", "args": [["Ligand"], 2.5], "kwargs": {}},
    })


def oracle_scores(batch, ind):
    tors, xf, rot = batch.pack([ind])
    eng = batch.engines[0][0]
    xyz = OA.express(eng.mobile, tors[0], xf[0], None if rot is None else rot[0])
    res = OA.contacts(eng.mobile, eng.receptor, eng.params, xyz)
    return res, xyz


def test_configs0_ga_run_matches_oracle(cuda):
    from gaudi_b200 import algorithms
    cfg = job(100, 10)
    pop, log, elite = algorithms.launch(cfg, seed=2026)
    assert len(pop) == 100 and len(log) == 11 and len(elite) >= 1
    assert all(ind.fitness.valid for ind in pop)
    batch = cfg._environment.batch
    assert len(batch.engines) == 1            # both Contacts objectives share one GPU context
    # the evolved population's fitness is what the oracle computes for the same alleles
    for ind in pop[::13]:
        res, xyz = oracle_scores(batch, ind)
        got = ind.fitness.values
        assert abs(got[0] - res.clash) <= 6e-4 + res.clash_slack + 1e-4 * abs(res.clash)
        assert abs(got[1] - res.hydro) <= 6e-4 + res.hydro_slack + 1e-4 * abs(res.hydro)
        lig = ind.find_molecule("Ligand")._expressed_coordinates
        assert np.abs(lig - xyz[batch.lig_slice]).max() < 1e-4
    # the GA did its job: the best clash score did not get worse than the random start
    random.seed(2026)
    first = [deep for deep in algorithms.launch(job(100, 0), seed=2026)[0]]
    assert min(i.fitness.values[0] for i in pop) <= min(i.fitness.values[0] for i in first)


def test_single_evaluate_and_similarity(cuda):
    from gaudi_b200 import algorithms, base, similarity
    cfg = job(12, 0)
    pop, _, _ = algorithms.launch(cfg, seed=5)
    env = cfg._environment
    batched = env.evaluate_population(pop)
    for ind, want in zip(pop[:4], batched):
        assert env.evaluate(ind) == want                      # population of one == row of the batch
        assert env.objectives["Clashes"].evaluate(ind) == want[0]
    sim = env.batch.similarity_matrix(pop, ["Ligand"], 2.5)
    ref = OA.rmsd_matrix(np.stack([i.find_molecule("Ligand")._expressed_coordinates for i in pop]))
    clear = np.abs(ref - 2.5) > 1e-4
    assert np.array_equal(sim[clear], (ref <= 2.5)[clear])
    assert pop[0].similar(pop[0]) and sim[0, 0]
    env.batch._sim = None                                     # no batch bitmask: pairwise GPU call
    for a, b in ((0, 1), (2, 7)):
        if clear[a, b]:
            assert similarity.rmsd(pop[a], pop[b], ["Ligand"], 2.5) == bool(ref[a, b] <= 2.5)


def test_rotamers_flexible_side_chains(cuda):
    """configs[4] in miniature: Rotamers gene on residues around the site + Torsion + Search."""
    from gaudi_b200 import base
    rng = np.random.default_rng(11)
    prot = G.load("5er1_protein_mol2")
    lig = G.load("5er1_ligand_mol2")
    # residues with side chains closest to the ligand
    d = np.linalg.norm(prot.xyz[:, None, :] - lig.xyz[None, ::4, :], axis=2).min(axis=1)
    near = []
    for a in np.argsort(d):
        r, nm = int(prot.resseq[a]), prot.resname[a][:3].upper()
        if nm not in ("ALA", "GLY", "PRO") and r not in near:
            near.append(r)
        if len(near) == 6:
            break
    library = {nm: rng.uniform(-180, 180, size=(int(rng.integers(3, 9)), 4)).tolist()
               for nm in set(prot.resname[a][:3].upper() for a in range(prot.n_atoms))}
    cfg = job(48, 0, rotamers=([("Protein", str(r)) for r in near], library))
    random.seed(3)
    pop = [base.MolecularIndividual(cfg) for _ in range(48)]
    env = base.Environment(cfg)
    cfg._environment = env
    scores = env.evaluate_population(pop)
    batch = env.batch
    assert batch.n_rot_res == 6 and batch.mobile.n_atoms > lig.n_atoms
    assert len(batch.receptor.ms_excl_idx) > 0        # side chains are bonded to the static backbone
    checked = 0
    for ind, got in list(zip(pop, scores))[::5]:
        res, _ = oracle_scores(batch, ind)
        assert abs(got[0] - res.clash) <= 6e-4 + res.clash_slack + 1e-4 * abs(res.clash), (got, res.clash)
        assert abs(got[1] - res.hydro) <= 6e-4 + res.hydro_slack + 1e-4 * abs(res.hydro), (got, res.hydro)
        checked += 1
    assert checked >= 9
    # express parity incl. the rotamer-driven torsions, and CA atoms never move (test_genes_rotamers.py:62-85)
    tors, xf, rot = batch.pack(pop[:8])
    eng = batch.engines[0][0]
    xyz = eng.express(tors, xf, rot).cpu().numpy().astype(np.float64) + eng.centre
    for p in range(8):
        assert np.abs(xyz[p] - OA.express(eng.mobile, tors[p], xf[p], rot[p])).max() < 1e-4
    fixed = np.unique(eng.mobile.tors_fixed[eng.mobile.tors_src < 0])
    ca_like = [int(a) for a in fixed
               if not any((int(m[int(a) >> 6]) >> (int(a) & 63)) & 1 for m in eng.mobile.tors_mask)]
    assert len(ca_like) >= 1
    assert np.abs(xyz[:, ca_like] - eng.mobile.xyz0[ca_like]).max() < 1e-5
    # Pose staging with flexible side chains: the fused kernel copies the 16-byte aligned span around
    # each pose (12 M bytes, here at every alignment) and reads the last pose of an array that does not
    # end on a 16-byte boundary with plain loads.  Warp-per-individual (2,701 and 2,700 poses) and
    # CTA-per-individual (49, 1): fused == tiled, bit for bit.
    import dataclasses
    from gaudi_b200.engine import Engine
    tiled = Engine(eng.mobile, eng.receptor, dataclasses.replace(eng.params, tiled=True), device=0)
    reps = -(-2701 // 48)
    T, X, R = (np.tile(v, (reps, 1))[:2701] for v in batch.pack(pop))
    T = T + rng.uniform(-20, 20, size=T.shape)
    for n in (2701, 2700, 49, 1):
        s1, c1, x1 = eng.evaluate(T[:n], X[:n], R[:n], want_xyz=True)
        s2, c2 = tiled.contacts(x1)
        assert cuda.equal(s1, s2) and cuda.equal(c1, c2), n
    tiled.close()


def test_contacts_without_same_residue_pairs(cuda):
    """Contacts(same_residue=False) = DetectClash intraRes=False (gaudi/objectives/contacts.py:142): pairs
    inside one residue — the whole ligand, a flexible side chain against its own backbone, static
    neighbours of one residue — are not reported.  Oracle, fused and tiled agree, and the scores differ
    from the default."""
    import dataclasses
    from gaudi_b200 import base
    from gaudi_b200.engine import Engine
    rng = np.random.default_rng(21)
    prot = G.load("5er1_protein_mol2")
    lig = G.load("5er1_ligand_mol2")
    d = np.linalg.norm(prot.xyz[:, None, :] - lig.xyz[None, ::4, :], axis=2).min(axis=1)
    near = []
    for a in np.argsort(d):
        r, nm = int(prot.resseq[a]), prot.resname[a][:3].upper()
        if nm not in ("ALA", "GLY", "PRO") and r not in near:
            near.append(r)
        if len(near) == 4:
            break
    library = {nm: rng.uniform(-180, 180, size=(5, 4)).tolist()
               for nm in set(prot.resname[a][:3].upper() for a in range(prot.n_atoms))}
    results = {}
    for same in (True, False):
        cfg = job(24, 0, rotamers=([("Protein", str(r)) for r in near], library))
        for o in cfg.objectives:
            o["same_residue"] = same
        random.seed(9)
        pop = [base.MolecularIndividual(cfg) for _ in range(24)]
        env = base.Environment(cfg)
        cfg._environment = env
        scores = np.array(env.evaluate_population(pop))
        batch = env.batch
        eng = batch.engines[0][0]
        assert eng.params.same_residue is same
        for ind, got in list(zip(pop, scores))[::4]:
            res, _ = oracle_scores(batch, ind)
            assert abs(got[0] - res.clash) <= 6e-4 + res.clash_slack + 1e-4 * abs(res.clash), (same, got, res.clash)
            assert abs(got[1] - res.hydro) <= 6e-4 + res.hydro_slack + 1e-4 * abs(res.hydro), (same, got, res.hydro)
        tiled = Engine(eng.mobile, eng.receptor, dataclasses.replace(eng.params, tiled=True), device=0)
        T, X, R = batch.pack(pop)
        s1, c1, x1 = eng.evaluate(T, X, R, want_xyz=True)
        s2, c2 = tiled.contacts(x1)
        assert cuda.equal(s1, s2) and cuda.equal(c1, c2)
        tiled.close()
        results[same] = scores
    assert not np.allclose(results[True], results[False])
    # the ligand is one residue: without same-residue pairs its internal contacts are gone
    assert (np.abs(results[False][:, 1]) <= np.abs(results[True][:, 1]) + 1e-6).all()


def test_many_flexible_side_chains(cuda):
    """36 flexible side chains: more than the 32 bodies the candidate expansion keeps masks for, so the
    non-seed slots fall back to plain runs of 32 (gmm_set_receptor); fused == tiled bit for bit in both
    the CTA-per-individual and the warp-per-individual regime, and the oracle agrees."""
    import dataclasses
    from gaudi_b200 import base
    from gaudi_b200.engine import Engine
    rng = np.random.default_rng(12)
    prot = G.load("5er1_protein_mol2")
    lig = G.load("5er1_ligand_mol2")
    d = np.linalg.norm(prot.xyz[:, None, :] - lig.xyz[None, ::4, :], axis=2).min(axis=1)
    near = []
    for a in np.argsort(d):
        r, nm = int(prot.resseq[a]), prot.resname[a][:3].upper()
        if nm not in ("ALA", "GLY", "PRO") and r not in near:
            near.append(r)
        if len(near) == 36:
            break
    library = {nm: rng.uniform(-180, 180, size=(int(rng.integers(3, 9)), 4)).tolist()
               for nm in set(prot.resname[a][:3].upper() for a in range(prot.n_atoms))}
    cfg = job(40, 0, rotamers=([("Protein", str(r)) for r in near], library))
    random.seed(5)
    pop = [base.MolecularIndividual(cfg) for _ in range(40)]
    env = base.Environment(cfg)
    cfg._environment = env
    scores = env.evaluate_population(pop)
    batch = env.batch
    eng = batch.engines[0][0]
    assert batch.n_rot_res == 36 and eng.mobile.n_atoms - lig.n_atoms > 32 * 8
    for ind, got in list(zip(pop, scores))[::8]:
        res, _ = oracle_scores(batch, ind)
        assert abs(got[0] - res.clash) <= 6e-4 + res.clash_slack + 1e-4 * abs(res.clash), (got, res.clash)
        assert abs(got[1] - res.hydro) <= 6e-4 + res.hydro_slack + 1e-4 * abs(res.hydro), (got, res.hydro)
    tiled = Engine(eng.mobile, eng.receptor, dataclasses.replace(eng.params, tiled=True), device=0)
    reps = -(-2700 // 40)
    T, X, R = (np.tile(v, (reps, 1))[:2700] for v in batch.pack(pop))
    T = T + rng.uniform(-20, 20, size=T.shape)
    for n in (2700, 40):
        s1, c1, x1 = eng.evaluate(T[:n], X[:n], R[:n], want_xyz=True)
        s2, c2 = tiled.contacts(x1)
        assert cuda.equal(s1, s2) and cuda.equal(c1, c2), n
    tiled.close()


def test_tensor_ga_configs0(cuda):
    """The array-population loop on the same job: fitness of the evolved population equals the
    oracle's for the same alleles, and a 2,000-individual population runs generations in a
    fraction of a second each."""
    from gaudi_b200 import tensor_ga
    cfg = job(100, 10)
    pop, log, batch = tensor_ga.launch(cfg, seed=7)
    assert pop["fitness"].shape == (100, 2) and len(log) == 11
    assert np.isfinite(pop["fitness"]).all()
    eng = batch.engines[0][0]
    xf = tensor_ga.TensorGA.xforms(pop)
    for p in range(0, 100, 17):
        xyz = OA.express(eng.mobile, pop["torsions"][p], xf[p])
        res = OA.contacts(eng.mobile, eng.receptor, eng.params, xyz)
        got = pop["fitness"][p]
        assert abs(got[0] - res.clash) <= 6e-4 + res.clash_slack + 1e-4 * abs(res.clash)
        assert abs(got[1] - res.hydro) <= 6e-4 + res.hydro_slack + 1e-4 * abs(res.hydro)
        assert np.abs(pop["coords"][p].astype(np.float64) + batch.centre - xyz[batch.lig_slice]).max() < 1e-4
    first = tensor_ga.launch(job(100, 0), seed=7)[0]
    assert pop["fitness"][:, 0].min() <= first["fitness"][:, 0].min()
    big, log, _ = tensor_ga.launch(job(2000, 3), seed=8)
    assert big["fitness"].shape == (2000, 2)
    assert all(entry["speed"] > 2000 for entry in log[1:]), log      # evaluations per second


def test_expressed_coordinates_of_every_molecule(cuda):
    """post_express (gaudi/base.py:301-303) caches coordinates of every Molecule gene, the protein
    included; with a Rotamers gene its flexible side chains carry the expressed positions."""
    from gaudi_b200 import base, similarity
    prot = G.load("5er1_protein_mol2")
    lig = G.load("5er1_ligand_mol2")
    library = {"LEU": [[60.0, 170.0], [-60.0, 60.0]], "ASP": [[-170.0, 20.0]], "TYR": [[180.0, 80.0]]}
    res = sorted({int(prot.resseq[a]) for a in range(prot.n_atoms)
                  if prot.resname[a][:3].upper() in library})[:5]
    cfg = job(6, 0, rotamers=([("Protein", str(r)) for r in res], library))
    random.seed(4)
    pop = [base.MolecularIndividual(cfg) for _ in range(6)]
    env = base.Environment(cfg)
    cfg._environment = env
    env.evaluate_population(pop)
    for ind in pop:
        p_xyz = ind.find_molecule("Protein")._expressed_coordinates
        l_xyz = ind.find_molecule("Ligand")._expressed_coordinates
        assert p_xyz.shape == (prot.n_atoms, 3) and l_xyz.shape == (lig.n_atoms, 3)
        moved = np.abs(p_xyz - prot.xyz).max(axis=1) > 1e-4
        assert moved.sum() <= 12 * len(res)               # only side-chain atoms of the chosen residues
        assert set(int(prot.resseq[a]) for a in np.flatnonzero(moved)) <= set(res)
    assert any((np.abs(i.find_molecule("Protein")._expressed_coordinates - prot.xyz) > 1e-4).any() for i in pop)
    assert similarity.rmsd(pop[0], pop[0], ["Ligand", "Protein"], 0.5)
    assert pop[0].similar(pop[0])


def test_reference_yaml_job_runs(cuda, tmp_path):
    """A YAML job with the reference's own module names and structure files on disk, run through
    ``algorithms.launch``: the drop-in the plugin API promises."""
    from gaudi_b200 import algorithms
    G.write_mol2(G.load("5er1_protein_mol2"), str(tmp_path / "protein.mol2"))
    G.write_mol2(G.load("5er1_ligand_mol2"), str(tmp_path / "ligand.mol2"))
    job = tmp_path / "job.yaml"
    job.write_text("""
output: {verbose: false, precision: 3}
ga: {population: 30, generations: 2, mu: 1.0, lambda_: 1.0}
similarity: {module: gaudi.similarity.rmsd, args: [[Ligand], 2.5], kwargs: {}}
genes:
  - {name: Protein, module: gaudi.genes.molecule, path: protein.mol2}
  - {name: Ligand, module: gaudi.genes.molecule, path: ligand.mol2}
  - {name: Torsion, module: gaudi.genes.torsion, target: Ligand, flexibility: 360}
  - {name: Search, module: gaudi.genes.search, target: Ligand/1, radius: 3.0, rotate: true}
objectives:
  - {name: Clashes, module: gaudi.objectives.contacts, which: clashes, probes: [Ligand], radius: 5.0, weight: -1.0}
  - {name: Contacts, module: gaudi.objectives.contacts, which: hydrophobic, probes: [Ligand], radius: 5.0, weight: -1.0}
""")
    cfg = parse.Settings(str(job))
    pop, log, elite = algorithms.launch(cfg, seed=3)
    assert len(pop) == 30 and len(log) == 3 and all(ind.fitness.valid for ind in pop)
    batch = cfg._environment.batch
    for ind in pop[::7]:
        res, xyz = oracle_scores(batch, ind)
        got = ind.fitness.values
        assert abs(got[0] - res.clash) <= 6e-4 + res.clash_slack + 1e-4 * abs(res.clash)
        assert abs(got[1] - res.hydro) <= 6e-4 + res.hydro_slack + 1e-4 * abs(res.hydro)


def test_run_writes_the_elite(cuda, tmp_path):
    """``algorithms.run`` = launch + dump_population (gaudi/cli/gaudi_run.py:137-178): the elite ends
    up on disk as the reference's zip / mol2 / yaml artefacts, and the ligand coordinates in them are
    the oracle's expression of the written alleles."""
    import zipfile
    import yaml
    from gaudi_b200 import algorithms, molio
    cfg = job(40, 3)
    cfg.output.path = str(tmp_path / "out")
    cfg.output.name = "dock"
    cfg.output.check_every = 2
    pop, log, elite, index = algorithms.run(cfg, seed=9)
    listing = yaml.safe_load(open(index))
    assert len(listing["GAUDI.results"]) == len(elite) >= 1
    assert (tmp_path / "out" / "check2" / "dock.gaudi-output").exists()          # the check_every dump
    batch = cfg._environment.batch
    for k, ind in enumerate(elite):
        zname = tmp_path / "out" / ("dock_%03d.zip" % k)
        assert listing["GAUDI.results"][zname.name] == list(ind.fitness.values)
        with zipfile.ZipFile(str(zname)) as z:
            z.extract("Ligand_%03d_Ligand.mol2" % k, str(tmp_path))
        back = molio.read_mol2(str(tmp_path / ("Ligand_%03d_Ligand.mol2" % k)))
        _, xyz = oracle_scores(batch, ind)
        assert np.abs(back.xyz - xyz[batch.lig_slice]).max() < 1e-4
