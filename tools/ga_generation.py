"""Wall time of one (mu + lambda) generation of the array GA (gaudi_b200/tensor_ga.py) at
BASELINE.json configs[3] scale: population 65,536 (131,072 candidates per niching pass), 40-atom /
10-torsion ligand, 20,000-atom receptor.  Prints one JSON line with the per-phase seconds.

    python tools/ga_generation.py [--pop 65536] [--gens 3]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pop", type=int, default=65536)
    ap.add_argument("--gens", type=int, default=3)
    ap.add_argument("--receptor", type=int, default=20000)
    a = ap.parse_args()
    from gaudi_b200 import synth, tensor_ga as TG
    from gaudi_b200.engine import Engine
    from gaudi_b200.model import ContactParams
    params = ContactParams()
    lig = synth.make_ligand(40, 10, seed=1)
    rec = synth.make_receptor(a.receptor, seed=2, params=params)
    eng = Engine(lig, rec, params)

    def evaluate(t, xf, r):
        scores, _, xyz = eng.evaluate_host(t, xf, None, want_xyz=True)
        return np.round(scores, 3), xyz

    def similarity(coords, threshold):
        _, bits = eng.rmsd_matrix(coords, threshold=threshold, values=False)
        return eng.similar_lists(bits, len(coords))

    spec = {"n_torsions": lig.n_allele_torsions, "flexibility": 360.0, "precision": 1, "cx_eta": 10.0,
            "mut_eta": 20.0, "mut_indpb": 0.1, "search_center": [0.0, 0.0, 0.0], "search_radius": 5.0,
            "rotate": True, "interpolation": 0.5, "n_rot_res": 0}
    ga = TG.TensorGA(spec, (-1.0, -1.0), evaluate, similarity, threshold=2.5, seed=1)
    pop = ga.random_population(a.pop)
    pop, log = ga.run(pop, a.pop, a.pop, 0.5, 0.5, a.gens)
    last = log[-1]["seconds"]
    print(json.dumps({"population": a.pop, "generations": a.gens,
                      "seconds_last_generation": {k: round(v, 4) for k, v in last.items()},
                      "all_generations": [round(g["seconds"]["generation"], 3) for g in log[1:]],
                      "best": [float(pop["fitness"][:, 0].min()), float(pop["fitness"][:, 1].min())]}))


if __name__ == "__main__":
    main()
