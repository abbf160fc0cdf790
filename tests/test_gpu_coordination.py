"""Coordination objective on the GPU (gmm_set_coordination / gmm_coordination) against the
reference's two golden scores (tests/test_objectives_coordination.py:33-60) and against the CPU
oracle (oracle/coordination_oracle.py) on seeded populations.  Scores are fp64 on both sides and
computed from the same expressed coordinates: the tolerance is 1e-7 (the point-set registration
runs up to ten EM steps; rounding differences between the Jacobi rotation used here and numpy's
SVD stay far below that)."""
import random

import numpy as np
import pytest

import golden_util as G
import oracle_adapter as OA
from gaudi_b200 import model as M
from gaudi_b200 import molio
from test_oracle_golden import coordination_site_4c3w

pytestmark = pytest.mark.gpu

TOL = 1e-7


@pytest.fixture(scope="module")
def cuda(built_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch


def complex_engine(mobile_residue=130, **site_kw):
    """4c3w with one residue (130 = the vanadium complex, 52 = the coordinating ASP) as the mobile
    part moved by a Search gene, everything else static."""
    from gaudi_b200.engine import Engine
    mol = G.load("4c3w_mol2")
    mobile = [i for i in range(mol.n_atoms) if int(mol.resseq[i]) == mobile_residue]
    _, site, g = coordination_site_4c3w(mobile=mobile, **site_kw)
    params = M.ContactParams(only_internal=True)
    radius = np.full(mol.n_atoms, 1.7)
    flags = np.zeros(mol.n_atoms, np.uint8)
    flags[mobile] = M.F_SEED | M.F_XFORM
    centre = mol.xyz[mobile].mean(axis=0)
    mob, rec, _ = M.split_molecule(mol.xyz, radius, flags, mol.bonds, mobile, params, origin_atom=mobile[0], centre=centre)
    eng = Engine(mob, rec, params, device=0)
    eng.set_coordination([site])
    return mol, site, g, eng, mob


def poses(mob, n, seed, spread=0.6, turn=0.35):
    """Search alleles near the crystal pose: small rotations about the origin atom and shifts."""
    rng = np.random.default_rng(seed)
    xf = np.zeros((n, 3, 4))
    for p in range(n):
        axis = rng.normal(size=3); axis /= np.linalg.norm(axis)
        ang = rng.normal(scale=turn) if p else 0.0
        K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
        xf[p, :, :3] = np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K.dot(K)
        xf[p, :, 3] = mob.xyz0[mob.origin_atom] + (rng.normal(scale=spread, size=3) if p else 0.0)
    return xf.reshape(n, 12)


def test_goldens_through_the_kernels(cuda):
    """The crystal pose (identity expression): 2.4696, and 3.2405 with both corrections."""
    for kw, key in (({}, "score"), ({"center_of_mass_correction": True, "distance_correction": True}, "score_corrected")):
        mol, site, g, eng, mob = complex_engine(**kw)
        xyz = eng.express(np.zeros((2, 0)), None)
        got = eng.coordination(xyz).cpu().numpy()
        assert got.shape == (2, 1)
        assert abs(got[0, 0] - g[key]) < g["tol"] and got[0, 0] == got[1, 0], got
        world = OA.coordination_world(site, xyz[0].cpu().numpy().astype(np.float64) + eng.centre)
        assert abs(got[0, 0] - OA.coordination(site, world)) < TOL
        eng.close()


@pytest.mark.parametrize("residue, kw, expect", [
    (130, {}, "score"),                                             # the metal moves with its complex
    (130, {"center_of_mass_correction": True, "distance_correction": True}, "score"),
    (52, {}, "score"),                                              # the metal is static, a coordinating ASP moves
    (130, {"prevent_intruders": False, "min_atoms": 8}, "any"),
    (130, {"distance": 0.5, "radius": 3.5}, "score"),
    (130, {"distance": 2.0}, "penalty"),      # |2.0 - d| > 1.0 fails for every bonded atom: the walk ends at once
    (52, {"template": "tetrahedral", "min_atoms": 5}, "both"),
])
def test_population_matches_oracle(cuda, residue, kw, expect):
    from gaudi_b200.objectives.coordination import GEOMETRIES
    kw = dict(kw)
    if "template" in kw:
        kw["template"] = np.array(GEOMETRIES[kw["template"]])
    mol, site, g, eng, mob = complex_engine(mobile_residue=residue, **kw)
    n = 300
    xf = poses(mob, n, seed=residue)
    xyz = eng.express(np.zeros((n, 0)), xf)
    got = eng.coordination(xyz).cpu().numpy()[:, 0]
    x64 = xyz.cpu().numpy().astype(np.float64) + eng.centre
    kinds = set()
    for p in range(0, n, 3):
        want = OA.coordination(site, OA.coordination_world(site, x64[p]))
        assert abs(got[p] - want) < TOL * max(1.0, abs(want)), (p, got[p], want)
        kinds.add("penalty" if want in (100.0, 200.0, 300.0, 400.0, 500.0, 1000.0) else "score")
    if expect == "score":
        assert "score" in kinds
    elif expect != "any":
        assert kinds == ({"score", "penalty"} if expect == "both" else {expect}), kinds
    eng.close()


def test_enforce_all_residues_is_the_penalty(cuda):
    """coordination.py:264 compares a set with a list: with enforce_all_residues the penalty is all
    the reference ever returns, and so does this."""
    mol, site, g, eng, mob = complex_engine(enforce_all_residues=True)
    xyz = eng.express(np.zeros((5, 0)), poses(mob, 5, seed=1))
    assert np.array_equal(eng.coordination(xyz).cpu().numpy()[:, 0], np.full(5, 1000.0))
    eng.close()


def test_plugin_path(cuda):
    """A Coordination objective in a job file next to Contacts: `module: gaudi.objectives.coordination`
    is redirected to the B200 plugin and the GA's fitness column equals the oracle's score."""
    import test_gpu_plugins as TP
    from gaudi_b200 import base, parse
    cfg = TP.job(40, 0)
    lig = G.load("5er1_ligand_mol2")
    probe = int(lig.serial[[i for i, e in enumerate(lig.element) if e.upper() == "N"][0]])
    data = dict(cfg)
    cfg = parse.Settings({
        "output": {"verbose": False, "precision": 6, "check_every": 0},
        "ga": {"population": 40, "generations": 0, "mu": 1.0, "lambda_": 1.0},
        "genes": [dict(g) for g in data["genes"]],
        "objectives": [dict(o) for o in data["objectives"]] + [
            {"name": "Site", "module": "gaudi.objectives.coordination", "probe": ["Ligand", probe],
             "residues": ["Protein/*"], "atom_elements": ["O", "N"], "radius": 4.5, "geometry": "tetrahedral",
             "min_atoms": 2, "prevent_intruders": False, "weight": -1.0}],
        "similarity": dict(data["similarity"]),
    })
    assert cfg.objectives[-1].module == "gaudi_b200.objectives.coordination"
    random.seed(4)
    pop = [base.MolecularIndividual(cfg) for _ in range(40)]
    env = base.Environment(cfg)
    cfg._environment = env
    scores = env.evaluate_population(pop)
    batch = env.batch
    site = batch.coordination[0][1].site(batch.system)
    eng = batch.engines[0][0]
    tors, xf, rot = batch.pack(pop)
    x64 = eng.express(tors, xf, rot).cpu().numpy().astype(np.float64) + eng.centre
    seen = set()
    for p in range(0, 40, 2):
        want = OA.coordination(site, OA.coordination_world(site, x64[p]))
        assert abs(scores[p][2] - round(want, 6)) < 2e-6, (p, scores[p], want)
        seen.add(want in (100.0, 200.0))
    assert False in seen
