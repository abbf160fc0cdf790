"""The reference's golden values reproduced ON THE GPU through the C ABI,
from the committed fixtures (nothing here reads /root/reference)."""
import numpy as np
import pytest

import golden_util as G
import oracle_adapter as OA
from gaudi_b200 import model as M
from gaudi_b200 import molio

pytestmark = pytest.mark.gpu
GV = G.values()


@pytest.fixture(scope="module")
def cuda(built_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch


def mobile_from(mol, rot, params, seed=True):
    return M.build_mobile(mol.xyz, molio.vdw_radii(mol), molio.atom_flags(mol, seed=seed, xform=True),
                          mol.bonds, [(f, m) for f, m, _ in rot], snap=[s for _, _, s in rot],
                          origin_atom=molio.nearest_atom_to_centre(mol), bond_separation=params.bond_separation)


def test_G1_backbone_torsion_on_gpu(cuda):
    from gaudi_b200.engine import Engine
    params = M.ContactParams(only_internal=True)
    mol = molio.perceive_bonds(G.load("1amb_pdb"))
    rot = molio.rotatable_bonds(mol, atom_types=(), atom_names=("CA",))
    mob = mobile_from(mol, rot, params)
    eng = Engine(mob, M.empty_receptor(mol.xyz.mean(axis=0)), params)
    x = eng.express(np.full((2, len(rot)), 90.0), None).cpu().numpy().astype(np.float64)
    g = GV["G1_backbone_torsion_distance"]
    assert abs(np.linalg.norm(x[0, 0] - x[0, -1]) - g["value"]) < g["tol"]
    ref = OA.express(mob, np.full(len(rot), 90.0), None)
    assert np.abs(x[1] + eng.centre - ref).max() < 1e-4


def test_G2_torsion_on_gpu(cuda):
    from gaudi_b200.engine import Engine
    params = M.ContactParams(only_internal=True)
    lig = G.ligand_3pk2()
    rot = molio.rotatable_bonds(lig)
    assert len(rot) == GV["G2_torsion"]["rotatable"]
    mob = mobile_from(lig, rot, params)
    assert mob.tors_snap.sum() == 1          # the C2=N2 bond only takes 0 / 180 (torsion.py:159-160)
    eng = Engine(mob, M.empty_receptor(lig.xyz.mean(axis=0)), params)
    x = eng.express(np.full((1, len(rot)), 1.0), None).cpu().numpy().astype(np.float64)[0]
    assert abs(np.linalg.norm(x[0] - x[-1]) - GV["G2_torsion"]["distance"]) < GV["G2_torsion"]["tol"]
    for allele in (np.full(len(rot), -77.3), np.linspace(-170, 170, len(rot))):
        got = eng.express(allele[None], None).cpu().numpy().astype(np.float64)[0] + eng.centre
        assert np.abs(got - OA.express(mob, allele, None)).max() < 1e-4


def test_G4_hydrophobic_1amb_on_gpu(cuda):
    """Whole peptide as the probe molecule: every pair is an intra-mobile pair."""
    from gaudi_b200.engine import Engine
    params = M.ContactParams(radius=7.3)
    mol = molio.perceive_bonds(G.load("1amb_pdb"))
    mob = mobile_from(mol, [], params)
    eng = Engine(mob, M.empty_receptor(mol.xyz.mean(axis=0)), params)
    # identity Search allele with origin 0: x' = R (x - origin) + t with t = origin
    origin = mol.xyz[mob.origin_atom]
    ident = np.tile(np.c_[np.eye(3), origin].reshape(1, 12), (3, 1))
    xyz = eng.express(np.zeros((3, 0)), ident)
    assert np.abs(xyz[0].cpu().numpy() + eng.centre - mol.xyz).max() < 1e-5
    s, c = eng.contacts(xyz)
    s = s.cpu().numpy()
    g = GV["G4_hydrophobic_1amb"]
    assert abs(s[0, 1] - g["value"]) < g["tol"], s
    assert int(c[0, 1]) == 31                # unique C-C pairs behind the score


def test_G3_clashes_5dfr_on_gpu(cuda):
    """2489-atom protein probing itself: first residues mobile, the rest static, all atoms zone
    seeds -> exercises mobile-static pairs, static-static pairs, exclusions across the cut."""
    from gaudi_b200.engine import Engine
    params = M.ContactParams(radius=4.2)
    mol = molio.perceive_bonds(G.load("5dfr_minimized_pdb"))
    rad, flags = molio.vdw_radii(mol), molio.atom_flags(mol, seed=True)
    for mobile_idx in (np.arange(0, 90), np.arange(1387, 1460)):   # N-terminus; around GLU 90
        mob, rec, _ = M.split_molecule(mol.xyz, rad, flags, mol.bonds, mobile_idx, params)
        eng = Engine(mob, rec, params)
        xyz = (mob.xyz0 - eng.centre).astype(np.float32)[None]
        s, c = eng.contacts(xyz)
        g = GV["G3_clashes_5dfr"]
        assert abs(float(s[0, 0]) - g["value"]) < g["tol"], s
        assert int(c[0, 0]) == 5             # five H...O- pairs
        res = OA.contacts(mob, rec, params, mob.xyz0)
        assert abs(float(s[0, 1]) - res.hydro) < 1e-4 * abs(res.hydro)


def test_G5_rmsd_on_gpu(cuda):
    from gaudi_b200.engine import Engine, bits_to_bool
    lig = G.ligand_3pk2()
    params = M.ContactParams(only_internal=True)
    eng = Engine(mobile_from(lig, [], params), M.empty_receptor(), params)
    a = G.load("3pk2_ligand_pdb").xyz
    b = G.load("3pk2_ligand_rotated_pdb").xyz
    centre = a.mean(axis=0)
    conf = np.stack([a, b, a]) - centre
    g = GV["G5_rmsd_3pk2"]
    out, bits = eng.rmsd_matrix(conf, threshold=g["value"])
    out = out.cpu().numpy()
    assert abs(out[0, 1] ** 2 - g["value"] ** 2) < g["tol_on_square"]
    assert out[0, 2] == 0.0
    sim = bits_to_bool(bits, 3)
    assert sim[0, 1] and sim[0, 2] and sim[1, 0]
    # recorded outputs of the reference's own gaudi/similarity.py
    z = np.load(G.GOLDEN + "/similarity_reference.npz")
    out, bits = eng.rmsd_matrix(z["conformers"], threshold=float(z["threshold"]))
    ref = np.sqrt(z["rmsd_squared"])
    assert np.abs(out.cpu().numpy() - ref).max() < 1e-5
    clear = np.abs(ref - float(z["threshold"])) > 1e-4
    assert np.array_equal(bits_to_bool(bits, ref.shape[0])[clear], z["similar"][clear])
