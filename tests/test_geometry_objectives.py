"""Distance and Angle objectives ("next" row f.4): the oracle restatement on hand-checkable
geometry (CPU), the C-ABI kernel against the oracle on random poses, and the plugin path (GPU)."""
import math

import numpy as np
import pytest

import golden_util as G
from gaudi_b200 import molio
import oracle_adapter as OA
from oracle import gaudi_oracle as O


def test_oracle_distance_cases():
    probes = np.array([[3.0, 0, 0], [0, 4.0, 0]])
    # distance.py:112-127: mean(|d - threshold|)
    assert O.distance_objective(probes, [0, 0, 0], 2.0) == pytest.approx((1.0 + 2.0) / 2)
    # a probe closer than threshold + tolerance is replaced by -1000 * weight
    assert O.distance_objective(probes, [0, 0, 0], 2.0, tolerance=1.5, weight=-1.0) == pytest.approx((1000.0 + 2.0) / 2)
    assert O.distance_objective(probes[:1], [3.0, 0, 0], 0.0) == 0.0


def test_oracle_angle_cases():
    a, b, c = [1.0, 0, 0], [0.0, 0, 0], [0, 1.0, 0]
    assert O.angle_objective([a, b, c], 90.0) == pytest.approx(0.0, abs=1e-12)
    assert O.angle_objective([a, b, c], 120.0) == pytest.approx(30.0)
    assert O.angle_objective([a, b, c], "planar") == pytest.approx(1.0)
    # dihedral (four points): trans = 180 -> planar 0; +60 gauche
    p = [[1.0, 0, 0], [0, 0, 0], [0, 0, 1.0], [-1.0, 0, 1.0]]
    assert O.angle_objective(p, "planar") == pytest.approx(0.0, abs=1e-12)
    assert abs(O.angle_objective(p, 0.0)) == pytest.approx(180.0)
    q = [[1.0, 0, 0], [0, 0, 0], [0, 0, 1.0], [math.cos(math.radians(60)), math.sin(math.radians(60)), 1.0]]
    assert abs(O.dihedral(*[np.array(x) for x in q])) == pytest.approx(60.0)


def test_distance_center_of_mass_oracle():
    pts = np.array([[1.0, 0, 0], [3.0, 0, 0]])
    assert O.distance_center_of_mass(pts, [1.0, 1.0], [2.0, 0, 0]) == 0.0
    assert O.distance_center_of_mass(pts, [3.0, 1.0], [0.0, 0, 0]) == pytest.approx(1.5)


@pytest.mark.gpu
def test_geometry_kernel_matches_oracle(built_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from gaudi_b200 import synth
    from gaudi_b200.engine import Engine
    from gaudi_b200.model import ContactParams
    params = ContactParams()
    mob, rec = synth.make_ligand(24, 6, seed=3), synth.make_receptor(400, seed=4, params=params)
    eng = Engine(mob, rec, params)
    P = 257
    tors, xf = synth.make_alleles(P, mob.n_allele_torsions, seed=8, radius=4.0)
    xyz = eng.express(tors, xf)
    fixed = np.asarray(rec.xyz[5], np.float64)
    lit = [1.5, -2.0, 0.25]
    terms = [
        {"kind": 0, "obj": 0, "points": [3, list(fixed)], "threshold": 2.5, "tolerance": -1.0, "weight": -1.0},
        {"kind": 0, "obj": 0, "points": [7, list(fixed)], "threshold": 2.5, "tolerance": -1.0, "weight": -1.0},
        {"kind": 0, "obj": 1, "points": [0, lit], "threshold": 4.0},
        {"kind": 1, "obj": 2, "points": [1, 2, list(fixed)], "threshold": 109.5},
        {"kind": 2, "obj": 3, "points": [0, 1, 2, 3], "threshold": 60.0},
        {"kind": 2, "obj": 4, "points": [list(fixed), 4, 5, 6], "threshold": "planar"},
        {"kind": 0, "obj": 5, "points": [9, 11], "threshold": 1.0},
        # Distance(center_of_mass=True): mass-weighted centre of three probes against a fixed target
        {"kind": 3, "obj": 6, "points": [2, list(fixed)], "threshold": 12.011},
        {"kind": 3, "obj": 6, "points": [8, list(fixed)], "threshold": 15.999},
        {"kind": 3, "obj": 6, "points": [13, list(fixed)], "threshold": 1.008},
    ]
    eng.set_geometry(terms, 7)
    got = eng.geometry(xyz).cpu().numpy()
    world = xyz.cpu().numpy().astype(np.float64) + eng.centre
    for p in range(P):
        x = world[p]
        want = [
            O.distance_objective(x[[3, 7]], fixed, 2.5, tolerance=-1.0, weight=-1.0),
            O.distance_objective(x[[0]], lit, 4.0),
            O.angle_objective([x[1], x[2], fixed], 109.5),
            O.angle_objective([x[0], x[1], x[2], x[3]], 60.0),
            O.angle_objective([fixed, x[4], x[5], x[6]], "planar"),
            O.distance_objective(x[[9]], x[11], 1.0),
            O.distance_center_of_mass(x[[2, 8, 13]], [12.011, 15.999, 1.008], fixed),
        ]
        # same fp32 coordinates on both sides: fp64 arithmetic agrees to rounding, except that
        # acos near 0/180 degrees amplifies it
        assert np.allclose(got[p], want, rtol=1e-9, atol=1e-6), (p, got[p], want)
    # the tolerance penalty did fire for some individual, and not for all
    pen = got[:, 0] >= 500.0
    assert 0 < pen.sum() < P or not pen.any()


@pytest.mark.gpu
def test_distance_and_angle_plugins(built_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from gaudi_b200 import algorithms, parse
    prot, lig = G.load("5er1_protein_mol2"), G.load("5er1_ligand_mol2")
    s_lig = [int(s) for s in lig.serial[:4]]
    s_prot = int(prot.serial[100])
    cfg = parse.Settings({
        "output": {"verbose": False, "precision": 3},
        "ga": {"population": 24, "generations": 1, "mu": 1.0, "lambda_": 1.0},
        "genes": [
            {"name": "Protein", "module": "gaudi_b200.genes.molecule", "path": "p.mol2", "mol": prot},
            {"name": "Ligand", "module": "gaudi_b200.genes.molecule", "path": "l.mol2", "mol": lig},
            {"name": "Torsion", "module": "gaudi_b200.genes.torsion", "target": "Ligand", "flexibility": 360},
            {"name": "Search", "module": "gaudi_b200.genes.search", "target": "Ligand/1", "radius": 3.0, "rotate": True},
        ],
        "objectives": [
            {"name": "Clashes", "module": "gaudi_b200.objectives.contacts", "which": "clashes",
             "probes": ["Ligand"], "radius": 5.0, "weight": -1.0},
            {"name": "Dist", "module": "gaudi_b200.objectives.distance", "probes": [["Ligand", s_lig[0]]],
             "target": ["Protein", s_prot], "threshold": 3.0, "weight": -1.0},
            {"name": "Point", "module": "gaudi_b200.objectives.distance",
             "probes": [["Ligand", s_lig[0]], ["Ligand", s_lig[1]]],
             "target": [float(v) for v in lig.xyz.mean(axis=0)], "threshold": 1.0, "tolerance": -0.5, "weight": -1.0},
            {"name": "Dihedral", "module": "gaudi_b200.objectives.angle",
             "probes": [["Ligand", s] for s in s_lig], "threshold": "planar", "weight": -1.0},
            {"name": "Centre", "module": "gaudi_b200.objectives.distance", "probes": [["Ligand", s] for s in s_lig[:3]],
             "target": ["Protein", s_prot], "threshold": 0.0, "center_of_mass": True, "weight": -1.0},
        ],
        "similarity": {"module": "gaudi_b200.similarity.rmsd", "args": [["Ligand"], 2.5], "kwargs": {}},
    })
    pop, _, _ = algorithms.launch(cfg, seed=3)
    batch = cfg._environment.batch
    rows = [int(np.flatnonzero(lig.serial == s)[0]) for s in s_lig]
    for ind in pop:
        x = ind.find_molecule("Ligand")._expressed_coordinates
        got = ind.fitness.values
        assert got[1] == pytest.approx(O.distance_objective(x[[rows[0]]], prot.xyz[100], 3.0), abs=2e-3)
        assert got[2] == pytest.approx(O.distance_objective(x[rows[:2]], lig.xyz.mean(axis=0), 1.0, tolerance=-0.5,
                                                            weight=-1.0), abs=2e-3)
        assert got[3] == pytest.approx(O.angle_objective([x[r] for r in rows], "planar"), abs=2e-3)
        masses = [molio.ATOMIC_MASS[lig.element[r].upper()] for r in rows[:3]]
        assert got[4] == pytest.approx(O.distance_center_of_mass(x[rows[:3]], masses, prot.xyz[100]), abs=2e-3)
    assert len(batch.engines) == 1


@pytest.mark.gpu
def test_three_objective_array_ga(built_lib):
    """Clashes + Contacts + Distance through the array GA: the restraint column comes from
    gmm_geometry, selection from the device non-dominated sort (three objectives)."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from gaudi_b200 import parse, tensor_ga
    prot, lig = G.load("5er1_protein_mol2"), G.load("5er1_ligand_mol2")
    s_lig, s_prot = int(lig.serial[0]), int(prot.serial[100])
    cfg = parse.Settings({
        "output": {"verbose": False, "precision": 3},
        "ga": {"population": 600, "generations": 2, "mu": 1.0, "lambda_": 1.0},
        "genes": [
            {"name": "Protein", "module": "gaudi_b200.genes.molecule", "path": "p.mol2", "mol": prot},
            {"name": "Ligand", "module": "gaudi_b200.genes.molecule", "path": "l.mol2", "mol": lig},
            {"name": "Torsion", "module": "gaudi_b200.genes.torsion", "target": "Ligand", "flexibility": 360},
            {"name": "Search", "module": "gaudi_b200.genes.search", "target": "Ligand/1", "radius": 3.0, "rotate": True},
        ],
        "objectives": [
            {"name": "Clashes", "module": "gaudi_b200.objectives.contacts", "which": "clashes",
             "probes": ["Ligand"], "radius": 5.0, "weight": -1.0},
            {"name": "Contacts", "module": "gaudi_b200.objectives.contacts", "which": "hydrophobic",
             "probes": ["Ligand"], "radius": 5.0, "weight": -1.0},
            {"name": "Dist", "module": "gaudi_b200.objectives.distance", "probes": [["Ligand", s_lig]],
             "target": ["Protein", s_prot], "threshold": 3.0, "weight": -1.0},
        ],
        "similarity": {"module": "gaudi_b200.similarity.rmsd", "args": [["Ligand"], 2.5], "kwargs": {}},
    })
    pop, log, batch = tensor_ga.launch(cfg, seed=11)
    fit = pop["fitness"]
    assert fit.shape == (600, 3) and np.isfinite(fit).all()
    row = int(np.flatnonzero(lig.serial == s_lig)[0])
    world = pop["coords"].astype(np.float64) + batch.centre
    for p in range(0, 600, 37):
        want = O.distance_objective(world[p, [row]], prot.xyz[100], 3.0)
        assert fit[p, 2] == pytest.approx(want, abs=2e-3)
    # the selected population starts with the non-dominated front of what it was selected from
    w = fit * np.array([-1.0, -1.0, -1.0])
    fronts = batch.engines[0][0].nondominated_fronts(w)
    assert (fronts == 0).sum() >= 1 and fronts.max() >= 1
    tensor_ga.set_fronts_backend(None)
