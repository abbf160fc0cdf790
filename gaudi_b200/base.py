"""Individual / Environment of the B200 path.

Mirrors gaudi/base.py: ``MolecularIndividual`` is an ordered container of gene plugins
(base.py:104-119,275-316), ``Environment`` an ordered container of objective plugins whose
``evaluate(individual)`` returns the rounded score list (base.py:342-364).  What is new is
``BatchEvaluator``: it flattens the genes of one template individual and the Contacts
objectives into the static model of ``gaudi_b200.model``, owns the GPU context(s), and scores a
whole list of individuals with one ``gmm_evaluate`` call — the body of the batched
``toolbox.map`` (gaudi/algorithms.py:99-102)."""
from __future__ import annotations

import sys
from collections import OrderedDict
from contextlib import contextmanager
from copy import deepcopy

import numpy as np

from . import model as M
from . import molio, plugin
from .deaplite import Fitness


class MolecularIndividual(object):
    _CACHE = {}

    def __init__(self, cfg=None, **kwargs):
        self.cfg = cfg
        self.genes = OrderedDict()
        self._molecules = OrderedDict()
        self.fitness = None
        self.expressed = False
        self.__ready__()
        self.__expression_hooks__()

    def __ready__(self):
        if self.cfg is not None:
            plugin.load_plugins(self.cfg.genes, container=self.genes, parent=self,
                                cx_eta=self.cfg.ga.cx_eta, mut_eta=self.cfg.ga.mut_eta,
                                mut_indpb=self.cfg.ga.mut_indpb)
            self.fitness = Fitness(self.cfg.weights)
            mod, fn = self.cfg.similarity.module.rsplit(".", 1)
            if mod not in sys.modules:
                __import__(mod)
            self._similarity = getattr(sys.modules[mod], fn)
        for name, gene in self.genes.items():
            if gene.__class__.__name__ == "Molecule":
                self._molecules[name] = gene
        for name, gene in self.genes.items():
            gene.__ready__()

    def __expression_hooks__(self):
        with expressed(self):
            for gene in self.genes.values():
                gene.__expression_hooks__()

    def __deepcopy__(self, memo):
        # gaudi/base.py:289-299 re-initialises and re-expresses every clone; here a clone is
        # a copy of the alleles (the "row copy" of SURVEY.md section 3.4).
        new = self.__class__.__new__(self.__class__)
        new.cfg = self.cfg
        new.genes = OrderedDict()
        new._molecules = OrderedDict()
        for name, gene in self.genes.items():
            if gene.__class__.__name__ == "Molecule":
                g = deepcopy(gene, memo)
            else:
                g = gene.__class__.__new__(gene.__class__)
                g.__dict__.update(gene.__dict__)
                g.allele = deepcopy(gene.allele)
            g.parent = new
            new.genes[name] = g
            if g.__class__.__name__ == "Molecule":
                new._molecules[name] = g
        new.fitness = deepcopy(self.fitness, memo)
        new._similarity = getattr(self, "_similarity", None)
        new.expressed = self.expressed
        return new

    def express(self):
        for gene in self.genes.values():
            gene.express()
        self.expressed = True

    def unexpress(self):
        for gene in reversed(list(self.genes.values())):
            gene.unexpress()
        self.expressed = False

    def mate(self, other):
        for a, b in zip(self.genes.values(), other.genes.values()):
            a.mate(b)
        return self, other

    def mutate(self, indpb):
        for gene in self.genes.values():
            gene.mutate(indpb)
        return self,

    def similar(self, other):
        return self._similarity(self, other, *self.cfg.similarity.args, **self.cfg.similarity.kwargs)

    def write(self, i, path=None):
        """gaudi/base.py:231-266: one zip per individual — every gene's file (Molecule genes: mol2
        with the expressed coordinates; the others: their allele as text) plus a ``.gaudi`` yaml
        naming them and the scores.  (The yaml is named after the LAST gene, as in the reference,
        whose loop variable shadows the job name.)"""
        import os
        import yaml
        from zipfile import ZipFile, ZIP_DEFLATED, ZIP_STORED
        path = path if path is not None else self.cfg.output.path
        name = self.cfg.output.name
        compress = ZIP_DEFLATED if self.cfg.output.compress else ZIP_STORED
        if any(g._expressed_coordinates is None for g in self._molecules.values()):
            env = getattr(self.cfg, "_environment", None)      # expression happens on the GPU, in batches
            if env is None:
                raise RuntimeError("write(): the individual has never been evaluated and the job has no environment")
            env.evaluate_population([self])
        zipfilename = os.path.join(path, "{}_{:03d}.zip".format(name, i))
        with ZipFile(zipfilename, "w", compress) as z:
            output = OrderedDict()
            for name, gene in self.genes.items():
                filename = gene.write(path, "{}_{:03d}".format(name, i))
                if filename:
                    z.write(filename, os.path.basename(filename))
                    os.remove(filename)
                    output[name] = os.path.basename(filename)
            output["score"] = [float(v) for v in self.fitness.values]
            z.writestr("{}_{:03d}.gaudi".format(name, i), yaml.safe_dump(dict(output), default_flow_style=False))
        return zipfilename

    def find_molecule(self, name):
        try:
            return self._molecules[name]
        except KeyError:
            raise KeyError("Molecule {} not found. Try with {}".format(name, sorted(self._molecules)))

    def xyz(self, gene=None):
        try:
            return self._molecules[gene]._expressed_coordinates
        except KeyError:
            return {name: mol._expressed_coordinates for name, mol in self._molecules.items()}

    def clear_cache(self):
        self._CACHE.clear()
        for gene in self.genes.values():
            gene.clear_cache()


@contextmanager
def expressed(*individuals):
    try:
        for individual in individuals:
            individual.express()
        yield individuals
    finally:
        for individual in individuals:
            individual.unexpress()


class Environment(object):
    """``device``: the GPU of this process.  Under a one-process-per-GPU launch (``torchrun``;
    ``torch.distributed`` initialised with more than one rank) the batched evaluate is sharded by
    individual over the ranks (``gaudi_b200.parallel.HostSharded``): the GA runs on rank 0, the other
    ranks sit in ``BatchEvaluator.serve()``; pass ``group=False`` to keep a process on its own."""

    def __init__(self, cfg=None, device=0, group=None, **kwargs):
        self.objectives = OrderedDict()
        self.cfg = cfg
        self.device = device
        self.group = group
        self.batch = None
        if cfg is not None:
            self.weights = cfg.weights
            plugin.load_plugins(cfg.objectives, container=self.objectives, zone=None, environment=self,
                                precision=cfg.output.precision)

    def attach(self, template):
        """Build the GPU model from one (template) individual; called once per job."""
        self.batch = BatchEvaluator(template, self, device=self.device, group=self.group)
        return self.batch

    def evaluate(self, individual):
        """gaudi/base.py:353-364 for one individual (a population of one on the GPU)."""
        return self.evaluate_population([individual])[0]

    def evaluate_population(self, individuals):
        if self.batch is None:
            self.attach(individuals[0])
        return self.batch.evaluate(individuals)

    def clear_cache(self):
        for obj in self.objectives.values():
            obj.clear_cache()


class BatchEvaluator(object):
    """Static model + GPU context(s) for one job, and the batched evaluate."""

    def __init__(self, template, environment, device=0, group=None):
        from .engine import Engine
        self.environment = environment
        self.device = device
        self.group = group
        self.sharded = None
        genes = template.genes
        self.torsion_name = next((n for n, g in genes.items() if g.__class__.__name__ == "Torsion"), None)
        self.search_name = next((n for n, g in genes.items() if g.__class__.__name__ == "Search"), None)
        self.rotamer_name = next((n for n, g in genes.items() if g.__class__.__name__ == "Rotamers"), None)
        names = list(genes.keys())
        tors, search, rot = (genes.get(self.torsion_name), genes.get(self.search_name),
                             genes.get(self.rotamer_name))
        if sum(g.__class__.__name__ == "Torsion" for g in genes.values()) > 1 or \
           sum(g.__class__.__name__ == "Search" for g in genes.values()) > 1:
            raise NotImplementedError("one Torsion and one Search gene per individual on the B200 path")
        self.ligand_name = tors.target if tors is not None else (search.target[0] if search is not None else None)
        if tors is not None and search is not None and tors.target != search.target[0]:
            raise NotImplementedError("Torsion and Search must target the same molecule on the B200 path")
        mol_names = list(template._molecules.keys())
        if self.ligand_name is None:
            self.ligand_name = mol_names[0]
        self.mol_names = mol_names

        # ---- one bonded system over every Molecule gene -------------------------------
        offsets, xyz, elements, bonds = {}, [], [], []
        radius, hb = [], []
        n = 0
        for name in mol_names:
            g = template._molecules[name]
            offsets[name] = n
            xyz.append(g.mol.xyz); elements += list(g.mol.element)
            bonds.append(np.asarray(g.mol.bonds, np.int64).reshape(-1, 2) + n)
            radius.append(g.radii()); hb.append(molio.hbond_flags(g.mol))
            n += g.mol.n_atoms
        self.offsets = offsets
        xyz = np.concatenate(xyz); radius = np.concatenate(radius); hb = np.concatenate(hb)
        bonds = np.concatenate(bonds) if bonds else np.zeros((0, 2), np.int64)
        lig = template._molecules[self.ligand_name]
        lo = offsets[self.ligand_name]
        lig_idx = np.arange(lo, lo + lig.mol.n_atoms)

        # rotatable bonds: ligand torsions first (allele columns), then rotamer chis
        rotatable, snap, tors_src = [], [], []
        if tors is not None:
            for k, (fx, mv, sn) in enumerate(tors.rotatable_bonds):
                rotatable.append((fx + lo, mv + lo)); snap.append(sn); tors_src.append(k)
        self.n_torsions = len(rotatable)
        flex_idx, cb_idx = [], []
        rot_ptr, rot_delta = [0], []
        if rot is not None:
            for r, (molname, pos, atoms, quads, chi0, table) in enumerate(rot.residues):
                ro = offsets[molname]
                g = template._molecules[molname]
                cb_idx += [a + ro for a in atoms if g.mol.name[a] == "CB"]
                rows = table if table else [list(chi0) if chi0 else [0.0]]
                for row in rows:
                    d = [(row[k] - chi0[k]) if (table and k < len(row) and k < len(chi0)) else 0.0 for k in range(4)]
                    rot_delta.append(d)
                rot_ptr.append(len(rot_delta))
                if not table:
                    continue
                adj = g.mol.adjacency()
                for k, q in enumerate(quads):
                    fx, mv = q[1], q[2]            # chi k turns about atoms 2-3, CA side fixed
                    side = M.side_of_bond(adj, fx, mv)
                    rotatable.append((fx + ro, mv + ro)); snap.append(0); tors_src.append(-(1 + 4 * r + k))
                    flex_idx += [a + ro for a in side] + [fx + ro]
        self.n_rot_res = len(rot_ptr) - 1 if rot is not None else 0
        mobile_idx = np.concatenate([lig_idx, np.array(sorted(set(flex_idx) - set(lig_idx.tolist())), np.int64)])
        self.mobile_idx = mobile_idx
        self.lig_slice = slice(0, lig.mol.n_atoms)
        # per molecule: (rows of the mobile array, atom indices inside the molecule) of its mobile atoms
        self._mobile_of = {}
        for name in mol_names:
            lo_m, n_m = offsets[name], template._molecules[name].mol.n_atoms
            rows = np.flatnonzero((mobile_idx >= lo_m) & (mobile_idx < lo_m + n_m))
            if len(rows):
                self._mobile_of[name] = (rows, mobile_idx[rows] - lo_m)

        origin = (search.origin_atom + lo) if search is not None else int(lig_idx[0])
        centre = np.asarray(search.center if search is not None else lig.mol.xyz.mean(axis=0), np.float64)
        t_before_s = True
        if tors is not None and search is not None:
            t_before_s = names.index(self.torsion_name) < names.index(self.search_name)

        # ---- one context per distinct Contacts parameter set ---------------------------
        self.groups = []            # [(engine, [(objective index, column)])]
        contacts = [(i, o) for i, o in enumerate(environment.objectives.values())
                    if o.__class__.__name__ == "Contacts"]
        self.geometry = [(i, o) for i, o in enumerate(environment.objectives.values())
                         if getattr(o, "geometry", False) is True]     # (Coordination.geometry is its template)
        self.coordination = [(i, o) for i, o in enumerate(environment.objectives.values())
                             if getattr(o, "coordination", False)]
        others = [o for o in environment.objectives.values()
                  if o.__class__.__name__ != "Contacts" and getattr(o, "geometry", False) is not True
                  and not getattr(o, "coordination", False)]
        if others:
            raise NotImplementedError("only Contacts, Distance, Angle and Coordination objectives run on the B200 "
                                      "path: %s" % [o.name for o in others])
        grouped = OrderedDict()
        for i, o in contacts:
            placed = False
            for key, g in grouped.items():
                if key[0] == o.model_key() and (g["penalty"] is None or o.penalty_key() in (None, g["penalty"])):
                    g["members"].append((i, o))
                    g["penalty"] = g["penalty"] or o.penalty_key()
                    placed = True
                    break
            if not placed:
                grouped[(o.model_key(), len(grouped))] = {"members": [(i, o)], "penalty": o.penalty_key()}
        self.engines = []
        if not grouped and (self.geometry or self.coordination):
            # restraints only: one context for the expression, with nothing in the Contacts zone
            grouped[(None, 0)] = {"members": [], "penalty": None}
        for key, g in grouped.items():
            o = g["members"][0][1] if g["members"] else None
            params = o.params() if o is not None else M.ContactParams(only_internal=True)
            if g["penalty"]:
                params.cutoff, params.weight = g["penalty"]
            else:
                params.cutoff = 0.0
            hyd = {e.upper() for e in o.hydrophobic_elements} if o is not None else set()
            flags = hb.copy()
            for a, e in enumerate(elements):
                if e in hyd:
                    flags[a] |= M.F_HYDROPHOBIC
            for pname in (o._probes if o is not None else [self.ligand_name]):
                po = offsets[pname]
                flags[po:po + template._molecules[pname].mol.n_atoms] |= M.F_SEED
            if o is not None and not o.only_internal:
                flags[np.asarray(cb_idx, np.int64)] |= M.F_SEED
            if search is not None:
                flags[lig_idx] |= M.F_XFORM
            mob, rec, _ = M.split_molecule(xyz, radius, flags, bonds, mobile_idx, params, rotatable=rotatable,
                                           origin_atom=origin, centre=centre,
                                           torsion_before_search=t_before_s)
            if not params.same_residue:
                res_id, seen = [], {}
                for mi, name in enumerate(mol_names):
                    mol = template._molecules[name].mol
                    res_id += [seen.setdefault((mi, int(r), str(ch)), len(seen)) for r, ch in zip(mol.resseq, mol.chain)]
                res_id = np.asarray(res_id)
                static_idx = np.setdiff1d(np.arange(len(xyz)), mobile_idx)
                M.exclude_same_residue(mob, rec, res_id[mobile_idx], res_id[static_idx])
            mob.tors_snap = np.asarray(snap, np.uint8)
            mob.tors_src = np.asarray(tors_src, np.int32)
            if rot is not None:
                mob.rot_ptr = np.asarray(rot_ptr, np.int32)
                mob.rot_delta = np.asarray(rot_delta, np.float64).reshape(-1, 4)
            eng = Engine(mob, rec, params, device=device)
            self.engines.append((eng, [(i, m.column) for i, m in g["members"]]))
        self.mobile, self.receptor = (self.engines[0][0].mobile, self.engines[0][0].receptor) if self.engines else (None, None)
        # ---- Distance / Angle restraints: one term table on the first context ------------
        if self.geometry:
            row_of = {int(a): r for r, a in enumerate(mobile_idx)}

            def locate(spec):
                molname, serial = spec
                mol = template._molecules[molname].mol
                hit = np.flatnonzero(mol.serial == int(serial))
                if len(hit) != 1:
                    raise ValueError("atom %s/%s not found" % (molname, serial))
                return offsets[molname] + int(hit[0])

            def resolve(spec):
                a = locate(spec)
                return row_of[a] if a in row_of else [float(v) for v in xyz[a]]

            def covalent(a, b):
                # chimera.Element.bondLength is not vendored: sum of the bond radii used by the
                # bond perception of molio (restated, unpinned)
                ea, eb = elements[locate(a)], elements[locate(b)]
                return molio.BOND_RADIUS.get(ea, 0.8) + molio.BOND_RADIUS.get(eb, 0.8)

            def mass(a):
                return molio.ATOMIC_MASS.get(elements[locate(a)], 12.011)

            terms = []
            for col, (i, o) in enumerate(self.geometry):
                kw = {"mass": mass} if getattr(o, "center_of_mass", False) else {}
                for t in o.terms(resolve, covalent, **kw):
                    terms.append(dict(t, obj=col))
            self.engines[0][0].set_geometry(terms, len(self.geometry))
        # ---- Coordination objectives: whole-job atom tables on the first context -------------
        if self.coordination:
            from types import SimpleNamespace
            row = -np.ones(len(xyz), np.int32)
            row[mobile_idx] = np.arange(len(mobile_idx), dtype=np.int32)
            neighbours = [[] for _ in range(len(xyz))]
            for a, b in bonds:                                  # atom.neighbors: in bond order
                neighbours[int(a)].append(int(b)); neighbours[int(b)].append(int(a))
            mol_of, resseq, atom_name, idatm = [], [], [], []
            for name in mol_names:
                mol = template._molecules[name].mol
                adj = mol.adjacency()
                mol_of += [name] * mol.n_atoms
                resseq += [int(r) for r in mol.resseq]
                atom_name += list(mol.name)
                idatm += [molio.idatm_type(mol, i, adj) for i in range(mol.n_atoms)]

            def locate_atom(spec):
                molname, serial = spec
                hit = np.flatnonzero(template._molecules[molname].mol.serial == int(serial))
                if len(hit) != 1:
                    raise ValueError("atom %s/%s not found" % (molname, serial))
                return offsets[molname] + int(hit[0])

            self.system = SimpleNamespace(xyz=np.asarray(xyz, np.float64), element=elements, idatm=idatm, name=atom_name,
                                          molecule=mol_of, resseq=resseq, neighbours=neighbours, mobile_row=row,
                                          locate=locate_atom)
            self.engines[0][0].set_coordination([o.site(self.system) for _, o in self.coordination])
        self.centre = centre
        self._sim = None

    # ---- allele packing ------------------------------------------------------------------
    def pack(self, individuals):
        P = len(individuals)
        tors = np.zeros((P, self.n_torsions))
        xf = np.zeros((P, 12)) if self.search_name else None
        rot = np.zeros((P, self.n_rot_res)) if self.n_rot_res else None
        for p, ind in enumerate(individuals):
            if self.torsion_name:
                a = ind.genes[self.torsion_name].allele
                tors[p, :min(len(a), self.n_torsions)] = a[:self.n_torsions]
            if self.search_name:
                xf[p] = ind.genes[self.search_name].device_allele()
            if rot is not None:
                rot[p] = ind.genes[self.rotamer_name].allele
        return tors, xf, rot

    # ---- scoring of allele arrays ----------------------------------------------------------
    def score_arrays(self, tors, xf, rot):
        """This process's GPU: raw objective columns ``(P, n_obj)`` fp64 (before the caller's
        ``round``) and device-frame coordinates ``(P, M, 3)`` float32 for allele arrays."""
        P = len(tors)
        objectives = list(self.environment.objectives.values())
        raw = np.zeros((P, len(objectives)))
        xyz = None
        for eng, members in self.engines:
            scores, _, coords = eng.evaluate_host(tors if self.n_torsions else np.zeros((P, 0)),
                                                  xf if self.search_name else None,
                                                  rot if self.n_rot_res else None, want_xyz=xyz is None)
            if coords is not None:
                xyz = coords
            for i, col in members:
                raw[:, i] = scores[:, col]
        if self.coordination:
            crd = self.engines[0][0].coordination(xyz).cpu().numpy()
            if np.isnan(crd).any():
                raise RuntimeError("Coordination: more than 192 atoms within the radius of a probe "
                                   "(gmm_coordination's per-individual limit); reduce `radius`")
            for col, (i, o) in enumerate(self.coordination):
                raw[:, i] = crd[:, col]
        if self.geometry:
            geo = self.engines[0][0].geometry(xyz).cpu().numpy()
            for col, (i, o) in enumerate(self.geometry):
                raw[:, i] = geo[:, col]
        return raw, xyz

    def enable_sharding(self, max_pop):
        """One-process-per-GPU launches: build the shared, page-locked population table (collective:
        every rank calls it with the same ``max_pop``).  No-op on a single process."""
        from . import parallel
        world, _ = parallel._world(self.group) if self.group is not False else (1, 0)
        if world == 1 or self.sharded is not None:
            return
        spec = OrderedDict()
        spec["tors"] = ((self.n_torsions,), np.float64)
        spec["xforms"] = ((12,), np.float64)
        spec["rot"] = ((self.n_rot_res,), np.float64)
        spec["raw"] = ((len(self.environment.objectives),), np.float64)
        spec["xyz"] = ((self.mobile.n_atoms, 3), np.float32)
        self.sharded = parallel.HostSharded(spec, max_pop, self.group)

    def _score_block(self, lo, hi):
        a = self.sharded.arrays
        raw, xyz = self.score_arrays(a["tors"][lo:hi], a["xforms"][lo:hi], a["rot"][lo:hi])
        a["raw"][lo:hi] = raw
        a["xyz"][lo:hi] = xyz

    def score_population(self, tors, xf, rot):
        """``score_arrays`` over every GPU of the job (rank 0 calls this, the others ``serve``)."""
        if self.sharded is None:
            return self.score_arrays(tors, xf, rot)
        a, P = self.sharded.arrays, len(tors)
        if P > self.sharded.max_rows:
            raise ValueError("population of %d exceeds the sharded table (%d rows)" % (P, self.sharded.max_rows))
        if self.n_torsions:
            a["tors"][:P] = tors
        if xf is not None:
            a["xforms"][:P] = xf
        if rot is not None and self.n_rot_res:
            a["rot"][:P] = rot
        self.sharded.run(P, self._score_block)
        return a["raw"][:P], a["xyz"][:P]

    def serve(self):
        """Ranks other than 0: evaluate blocks of rank 0's populations until it shuts the job down."""
        self.sharded.serve(self._score_block)

    def shutdown(self):
        if self.sharded is not None:
            self.sharded.shutdown()

    def evaluate(self, individuals):
        """Batched ``Environment.evaluate``: list of rounded score lists, in objective order
        (``round(score, obj.precision)``, gaudi/base.py:361)."""
        if not individuals:
            return []
        tors, xf, rot = self.pack(individuals)
        objectives = list(self.environment.objectives.values())
        raw, xyz = self.score_population(tors, xf, rot)
        out = [[round(float(raw[p, i]), objectives[i].precision) for i in range(len(objectives))]
               for p in range(len(individuals))]
        # post_express (gaudi/base.py:301-303): every Molecule gene gets its expressed, world-frame
        # coordinates.  Molecules without mobile atoms share one array (the reference copies the
        # whole protein on every evaluation); molecules with flexible side chains get a copy with
        # the moved atoms written in.
        world = xyz.astype(np.float64) + self.centre
        for p, ind in enumerate(individuals):
            for name, gene in ind._molecules.items():
                sel = self._mobile_of.get(name)
                if sel is None:
                    gene._expressed_coordinates = gene.mol.xyz
                elif name == self.ligand_name and len(sel[0]) == gene.mol.n_atoms:
                    gene._expressed_coordinates = world[p, self.lig_slice]
                else:
                    full = gene.mol.xyz.copy()
                    full[sel[1]] = world[p, sel[0]]
                    gene._expressed_coordinates = full
        return out

    def score_of(self, individual, objective):
        idx = list(self.environment.objectives.values()).index(objective)
        return self.evaluate([individual])[0][idx]

    # ---- batched similarity -----------------------------------------------------------------
    def _similarity_bits(self, individuals, subjects, threshold):
        """Device bitmask (N, ceil(N/32)) int32: pairs similar in EVERY subject molecule."""
        eng = self.engines[0][0]
        bits = None
        for s in subjects:
            conf = np.stack([np.asarray(ind._molecules[s]._expressed_coordinates - self.centre, np.float32)
                             for ind in individuals])
            _, b = eng.rmsd_matrix(conf, threshold=threshold, values=False)
            bits = b if bits is None else bits & b
        return bits

    def similarity_matrix(self, individuals, subjects, threshold):
        """One ``gmm_rmsd_matrix`` call per subject for every pair among ``individuals``; afterwards
        ``gaudi_b200.similarity.rmsd`` answers from the result.  Returns the (N, N) bool matrix."""
        from .engine import bits_to_bool
        bits = self._similarity_bits(individuals, subjects, threshold)
        ptr, idx = self.engines[0][0].similar_lists(bits, len(individuals))
        self._remember(individuals, ptr, idx, subjects, threshold)
        return bits_to_bool(bits, len(individuals))

    def similarity_lists(self, individuals, subjects, threshold):
        """As ``similarity_matrix`` but only the partner lists leave the device: ``(ptr, idx)`` CSR
        over ``individuals`` (self excluded), what the niching pass consumes."""
        bits = self._similarity_bits(individuals, subjects, threshold)
        ptr, idx = self.engines[0][0].similar_lists(bits, len(individuals))
        self._remember(individuals, ptr, idx, subjects, threshold)
        return ptr, idx

    def _remember(self, individuals, ptr, idx, subjects, threshold):
        # keyed by identity: the entry keeps the individuals alive so that no later individual can
        # be handed the id of one of these and read its row by mistake
        self._sim = ({id(ind): k for k, ind in enumerate(individuals)}, ptr, idx, tuple(subjects), float(threshold),
                     list(individuals))

    def lookup_similarity(self, ind1, ind2, subjects, threshold):
        if self._sim is None:
            return None
        index, ptr, idx, subj, thr, _alive = self._sim
        a, b = index.get(id(ind1)), index.get(id(ind2))
        if a is None or b is None or subj != tuple(subjects) or thr != float(threshold):
            return None
        if a == b:
            return True
        row = idx[ptr[a]:ptr[a + 1]]                       # partners of a, ascending
        k = int(np.searchsorted(row, b))
        return bool(k < len(row) and row[k] == b)

    def close(self):
        if self.sharded is not None:
            self.sharded.close()
            self.sharded = None
        for eng, _ in self.engines:
            eng.close()
