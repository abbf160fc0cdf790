"""ctypes face of oracle/gaudi_oracle.c (TEST INFRASTRUCTURE — see that file's header).

``COracle(mobile, receptor, params)`` takes the product's static model objects (read as plain
arrays) and evaluates whole populations on the host cores with OpenMP."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libgaudi_oracle.so")


class _Params(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("zone_radius", "clash_thr", "hydro_thr", "hb_allow", "amv", "cutoff", "weight")] + \
               [("only_internal", C.c_int)]


def _lib():
    if not os.path.exists(LIB):
        import subprocess
        subprocess.check_call(["make", "-s", "-C", HERE])
    lib = C.CDLL(LIB)
    lib.go_grid_build.restype = C.c_void_p
    lib.go_grid_build.argtypes = [C.c_int, C.c_void_p, C.c_double]
    lib.go_grid_free.argtypes = [C.c_void_p]
    lib.go_num_threads.restype = C.c_int
    return lib


def _within(adj, src, depth):
    seen = {src}
    frontier = [src]
    for _ in range(depth):
        nxt = []
        for a in frontier:
            for b in adj[a]:
                if b not in seen:
                    seen.add(b)
                    nxt.append(b)
        frontier = nxt
    return seen


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class COracle(object):
    def __init__(self, mob, rec, params):
        self.lib = _lib()
        self.mob, self.rec, self.params = mob, rec, params
        m, r = mob.n_atoms, rec.n_atoms
        self.M, self.R = m, r
        self.xyz0 = np.ascontiguousarray(mob.xyz0, np.float64)
        self.sta = np.ascontiguousarray(rec.xyz, np.float64).reshape(-1, 3)
        # radii as the device sees them (fp32-rounded), like tests/oracle_adapter.py
        self.radius = np.concatenate([np.asarray(mob.radius, np.float32), np.asarray(rec.radius, np.float32)]).astype(np.float64)
        self.flags = np.ascontiguousarray(np.concatenate([mob.flags, rec.flags]), np.uint8)
        # bond graph over all atoms -> "within <= bond_separation bonds" lists (DetectClash exclusions)
        adj = [[] for _ in range(m + r)]
        bonds = [(int(a), int(b)) for a, b in mob.bonds] + [(int(a) + m, int(b) + m) for a, b in rec.bonds]
        extra = getattr(rec, "ms_bonds", None)
        if extra is not None:
            bonds += [(int(a), int(b) + m) for a, b in extra]
        for a, b in bonds:
            adj[a].append(b); adj[b].append(a)
        ptr, idx = [0], []
        for a in range(m + r):
            if adj[a]:
                idx.extend(sorted(_within(adj, a, params.bond_separation)))
            ptr.append(len(idx))
        self.excl_ptr = np.asarray(ptr, np.int32)
        self.excl_idx = np.asarray(idx if idx else [0], np.int32)
        # torsions
        self.fixed = np.ascontiguousarray(mob.tors_fixed, np.int32)
        self.moving = np.ascontiguousarray(mob.tors_moving, np.int32)
        mptr, midx = [0], []
        for t in range(mob.n_torsions):
            midx.extend(int(i) for i in mob.moving_indices(t))
            mptr.append(len(midx))
        self.mask_ptr = np.asarray(mptr, np.int32)
        self.mask_idx = np.asarray(midx if midx else [0], np.int32)
        self.snap = np.ascontiguousarray(mob.tors_snap, np.uint8)
        self.tors_src = np.ascontiguousarray(mob.tors_src, np.int32)
        self.rot_ptr = np.ascontiguousarray(mob.rot_ptr, np.int32)
        self.rot_delta = np.ascontiguousarray(mob.rot_delta, np.float64).reshape(-1, 4)
        if self.rot_delta.shape[0] == 0:
            self.rot_delta = np.zeros((1, 4))
        cell = max(2.5, min(6.0, float(params.radius) if params.radius > 0 else 4.0))
        self.grid = self.lib.go_grid_build(r, _p(self.sta), C.c_double(cell))
        self.prm = _Params(params.radius, params.clash_threshold, params.hydrophobic_threshold,
                           params.hbond_allowance, params.assumed_max_vdw, params.cutoff, params.weight,
                           int(params.only_internal))

    def __del__(self):
        try:
            self.lib.go_grid_free(C.c_void_p(self.grid))
        except Exception:
            pass

    @property
    def threads(self):
        return int(self.lib.go_num_threads())

    def evaluate(self, tors, xforms, rot=None, want_xyz=False):
        tors = np.ascontiguousarray(tors if tors is not None else np.zeros((len(xforms), 0)), np.float64)
        P = tors.shape[0]
        xf = None if xforms is None else np.ascontiguousarray(xforms, np.float64)
        rt = None if rot is None else np.ascontiguousarray(rot, np.float64)
        scores = np.empty((P, 2))
        counts = np.empty((P, 2), np.int32)
        xyz = np.empty((P, self.M, 3)) if want_xyz else None
        def run(lo, hi):
            sl = slice(lo, hi)
            self.lib.go_evaluate_population(
                C.c_int(hi - lo), C.c_int(self.M), _p(self.xyz0), C.c_int(self.mob.n_torsions), _p(self.fixed),
                _p(self.moving), _p(self.mask_ptr), _p(self.mask_idx), _p(self.snap), _p(self.tors_src),
                C.c_int(tors.shape[1]), _p(tors[sl]), _p(None if xf is None else xf[sl]),
                C.c_int(self.mob.origin_atom), C.c_int(int(self.mob.torsion_before_search)),
                C.c_int(self.mob.n_rotamer_residues), _p(self.rot_ptr), _p(self.rot_delta),
                _p(None if rt is None else rt[sl]), C.c_int(self.R), _p(self.sta), C.c_void_p(self.grid),
                _p(self.radius), _p(self.flags), _p(self.excl_ptr), _p(self.excl_idx), C.byref(self.prm),
                _p(scores[sl]), _p(counts[sl]), _p(None if xyz is None else xyz[sl]))
        cores = os.cpu_count() or 1
        if self.threads > 1 or cores == 1 or P < 2 * cores:
            run(0, P)                      # OpenMP inside (or nothing to split)
        else:                              # no OpenMP in this build: slices on Python threads
            from concurrent.futures import ThreadPoolExecutor
            b = [(P * k) // (4 * cores) for k in range(4 * cores + 1)]
            with ThreadPoolExecutor(cores) as ex:
                list(ex.map(lambda k: run(b[k], b[k + 1]), range(4 * cores)))
        return scores, counts, xyz

    @property
    def workers(self):
        """Host threads evaluate() actually uses."""
        return self.threads if self.threads > 1 else (os.cpu_count() or 1)

    def contacts(self, xyz_mobile):
        mob = np.ascontiguousarray(xyz_mobile, np.float64)
        scores = np.empty(2)
        counts = np.empty(2, np.int32)
        self.lib.go_contacts(C.c_int(self.M), _p(mob), C.c_int(self.R), _p(self.sta), C.c_void_p(self.grid),
                             _p(self.radius), _p(self.flags), _p(self.excl_ptr), _p(self.excl_idx),
                             C.byref(self.prm), _p(scores), _p(counts))
        return scores, counts

    def rmsd_matrix(self, xyz, row0=0, row1=None):
        xyz = np.ascontiguousarray(xyz, np.float64)
        n, a = xyz.shape[0], xyz.shape[1]
        row1 = n if row1 is None else row1
        out = np.empty((row1 - row0, n))
        self.lib.go_rmsd_matrix(C.c_int(n), C.c_int(a), _p(xyz), C.c_int(row0), C.c_int(row1), _p(out))
        return out
