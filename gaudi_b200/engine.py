"""Python face of the C ABI: one ``Engine`` = one ``gmm_ctx`` (one GPU, one static model).

PyTorch is used only as the owner of device buffers and streams; every computation is a
call into ``libgaudimm_b200.so`` with raw ``data_ptr()`` device pointers on torch's current
stream.  There is no CPU path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .model import ContactParams, MobileModel, ReceptorModel


def _np_ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _dev_ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class Engine(object):
    """Batched express + Contacts + RMSD on one B200.

    Mirrors, for a whole population, what ``Environment.evaluate`` (gaudi/base.py:353-364)
    does for one individual and what ``gaudi.similarity.rmsd`` (gaudi/similarity.py:36-70)
    does for one pair.
    """

    def __init__(self, mobile: MobileModel, receptor: ReceptorModel,
                 params: ContactParams | None = None, device: int = 0):
        if not torch.cuda.is_available():
            raise RuntimeError("gaudi_b200 needs a CUDA device (there is no CPU fallback)")
        self._L = _lib.lib()
        self.device = int(device)
        self.torch_device = torch.device("cuda", self.device)
        self.mobile, self.receptor = mobile, receptor
        self.params = params or ContactParams()
        h = C.c_void_p()
        _lib.check(self._L.gmm_create(C.byref(h), self.device))
        self._h = h
        p = self.params
        opts = ((_lib.OPT_ONLY_INTERNAL if p.only_internal else 0) | (0 if p.cull else _lib.OPT_NO_CULL)
                | (_lib.OPT_TILED if p.tiled else 0)
                | (_lib.OPT_SEQUENTIAL_EXPRESS if getattr(p, 'sequential_express', False) else 0)
                | (_lib.OPT_RMSD_SIMT if getattr(p, 'rmsd_simt', False) else 0)
                | (_lib.OPT_RMSD_HMMA if getattr(p, 'rmsd_hmma', False) else 0))
        _lib.check(self._L.gmm_set_params(h, p.radius, p.clash_threshold, p.hydrophobic_threshold,
                                          p.hbond_allowance, p.assumed_max_vdw, p.cutoff, p.weight,
                                          opts))
        m = mobile
        keep = [np.ascontiguousarray(m.xyz0, np.float64), np.ascontiguousarray(m.radius, np.float32),
                np.ascontiguousarray(m.flags, np.uint8), np.ascontiguousarray(m.tors_fixed, np.int32),
                np.ascontiguousarray(m.tors_moving, np.int32), np.ascontiguousarray(m.tors_mask, np.uint64),
                np.ascontiguousarray(m.tors_snap, np.uint8), np.ascontiguousarray(m.tors_src, np.int32),
                np.ascontiguousarray(m.excl_mask, np.uint64)]
        _lib.check(self._L.gmm_set_mobile(h, m.n_atoms, _np_ptr(keep[0]), _np_ptr(keep[1]), _np_ptr(keep[2]),
                                          m.n_torsions, _np_ptr(keep[3]), _np_ptr(keep[4]), _np_ptr(keep[5]),
                                          _np_ptr(keep[6]), _np_ptr(keep[7]), _np_ptr(keep[8]),
                                          m.origin_atom, int(m.torsion_before_search)))
        if m.n_rotamer_residues:
            rp = np.ascontiguousarray(m.rot_ptr, np.int32)
            rd = np.ascontiguousarray(m.rot_delta, np.float64)
            _lib.check(self._L.gmm_set_rotamers(h, m.n_rotamer_residues, _np_ptr(rp), _np_ptr(rd)))
        r = receptor
        rk = [np.ascontiguousarray(r.xyz, np.float64), np.ascontiguousarray(r.radius, np.float32),
              np.ascontiguousarray(r.flags, np.uint8), np.ascontiguousarray(r.rr_ptr, np.int32),
              np.ascontiguousarray(r.rr_idx, np.int32), np.ascontiguousarray(r.centre, np.float64)]
        ms_ptr = ms_idx = None
        if len(r.ms_excl_idx):
            ms_ptr = np.ascontiguousarray(r.ms_excl_ptr, np.int32)
            ms_idx = np.ascontiguousarray(r.ms_excl_idx, np.int32)
        _lib.check(self._L.gmm_set_receptor(h, r.n_atoms, _np_ptr(rk[0]), _np_ptr(rk[1]), _np_ptr(rk[2]),
                                            _np_ptr(rk[3]), _np_ptr(rk[4]) if len(rk[4]) else None,
                                            _np_ptr(ms_ptr), _np_ptr(ms_idx), _np_ptr(rk[5])))
        self.centre = np.array(r.centre, dtype=np.float64)

    # -- plumbing ----------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._L.gmm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.torch_device).cuda_stream)

    def _dev(self, a, dtype):
        if a is None:
            return None
        if isinstance(a, torch.Tensor):
            return a.to(device=self.torch_device, dtype=dtype).contiguous()
        return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(self.torch_device)

    def reserve(self, max_pop):
        _lib.check(self._L.gmm_reserve(self._h, int(max_pop)))

    def reserve_rmsd(self, max_n, atoms):
        """Pre-size the workspace of bitmask-only ``rmsd_matrix`` calls (``gmm_reserve_rmsd``)."""
        _lib.check(self._L.gmm_reserve_rmsd(self._h, int(max_n), int(atoms)))

    # -- hot path ----------------------------------------------------------------------
    def express(self, torsions=None, xforms=None, rot_alleles=None, out=None):
        """(P, M, 3) float32 device-frame coordinates (add ``self.centre`` for world frame)."""
        t = self._dev(torsions, torch.float64)
        x = self._dev(xforms, torch.float64)
        r = self._dev(rot_alleles, torch.float64)
        P = int((t if t is not None else (x if x is not None else r)).shape[0])
        if out is None:
            out = torch.empty((P, self.mobile.n_atoms, 3), dtype=torch.float32, device=self.torch_device)
        _lib.check(self._L.gmm_express(self._h, P, _dev_ptr(t), _dev_ptr(x), _dev_ptr(r), _dev_ptr(out),
                                       self._stream()))
        return out

    def contacts(self, xyz, counts=True):
        xyz = self._dev(xyz, torch.float32)
        P = int(xyz.shape[0])
        scores = torch.empty((P, 2), dtype=torch.float64, device=self.torch_device)
        cnt = torch.empty((P, 2), dtype=torch.int32, device=self.torch_device) if counts else None
        _lib.check(self._L.gmm_contacts(self._h, P, _dev_ptr(xyz), _dev_ptr(scores), _dev_ptr(cnt),
                                        self._stream()))
        return scores, cnt

    def evaluate(self, torsions=None, xforms=None, rot_alleles=None, want_xyz=False, counts=True,
                 scores_out=None, counts_out=None, xyz_out=None):
        """Device-resident batched ``Environment.evaluate``: returns (scores, counts, xyz)."""
        t = self._dev(torsions, torch.float64)
        x = self._dev(xforms, torch.float64)
        r = self._dev(rot_alleles, torch.float64)
        P = int((t if t is not None else (x if x is not None else r)).shape[0])
        scores = scores_out if scores_out is not None else torch.empty(
            (P, 2), dtype=torch.float64, device=self.torch_device)
        cnt = counts_out
        if cnt is None and counts:
            cnt = torch.empty((P, 2), dtype=torch.int32, device=self.torch_device)
        xyz = xyz_out
        if xyz is None and want_xyz:
            xyz = torch.empty((P, self.mobile.n_atoms, 3), dtype=torch.float32, device=self.torch_device)
        _lib.check(self._L.gmm_evaluate(self._h, P, _dev_ptr(t), _dev_ptr(x), _dev_ptr(r), _dev_ptr(xyz),
                                        _dev_ptr(scores), _dev_ptr(cnt), self._stream()))
        return scores, cnt, xyz

    def evaluate_host(self, torsions=None, xforms=None, rot_alleles=None, want_xyz=False,
                      scores_out=None, counts_out=None, xyz_out=None):
        """Host buffers in, host buffers out (H2D + kernels + D2H inside the call).

        Inputs may be numpy arrays or CPU torch tensors; page-locked ones
        (``torch.empty(..., pin_memory=True)``) are copied from directly, pageable ones are
        staged.  ``scores_out`` / ``counts_out`` / ``xyz_out``: optional preallocated (pinned) CPU
        tensors or numpy arrays (for instance rows of a ``parallel.HostSharded`` table)."""
        def host(a):
            if a is None:
                return None, None
            if isinstance(a, torch.Tensor):
                a = a if (a.dtype == torch.float64 and a.is_contiguous()) else a.to(torch.float64).contiguous()
                return a, C.c_void_p(a.data_ptr())
            a = np.ascontiguousarray(a, np.float64)
            return a, _np_ptr(a)
        t, tp = host(torsions)
        x, xp = host(xforms)
        r, rp = host(rot_alleles)
        P = int((t if t is not None else (x if x is not None else r)).shape[0])
        if scores_out is None:
            scores_out = np.empty((P, 2), np.float64)
        if counts_out is None:
            counts_out = np.empty((P, 2), np.int32)
        sp = C.c_void_p(scores_out.data_ptr()) if isinstance(scores_out, torch.Tensor) else _np_ptr(scores_out)
        cp = C.c_void_p(counts_out.data_ptr()) if isinstance(counts_out, torch.Tensor) else _np_ptr(counts_out)
        xyz = xyz_out if xyz_out is not None else (np.empty((P, self.mobile.n_atoms, 3), np.float32) if want_xyz else None)
        xyzp = C.c_void_p(xyz.data_ptr()) if isinstance(xyz, torch.Tensor) else _np_ptr(xyz)
        _lib.check(self._L.gmm_evaluate_host(self._h, P, tp, xp, rp, xyzp, sp, cp, self._stream()))
        return scores_out, counts_out, xyz

    def rmsd_matrix(self, xyz, row0=0, row1=None, threshold=2.5, values=True, bits=True):
        """Rows [row0, row1) of the all-pairs RMSD matrix of (N, A, 3) conformers.
        Returns (rmsd float32 (rows, N) or None, similar uint32 bitmask (rows, ceil(N/32)) or None)."""
        xyz = self._dev(xyz, torch.float32)
        n, a = int(xyz.shape[0]), int(xyz.shape[1])
        row1 = n if row1 is None else int(row1)
        rows = row1 - row0
        out = torch.empty((rows, n), dtype=torch.float32, device=self.torch_device) if values else None
        # int32 storage for the uint32 words (torch has no general uint32 support)
        bm = torch.empty((rows, (n + 31) // 32), dtype=torch.int32, device=self.torch_device) if bits else None
        _lib.check(self._L.gmm_rmsd_matrix(self._h, n, a, _dev_ptr(xyz), int(row0), row1, _dev_ptr(out),
                                           _dev_ptr(bm), float(threshold), self._stream()))
        return out, bm

    def express_is_tree(self):
        """True when this model is expressed in the kinematic-tree form (see ``gmm_express_is_tree``)."""
        rc = self._L.gmm_express_is_tree(self._h)
        if rc < 0:
            _lib.check(rc)
        return bool(rc)

    def nondominated_fronts(self, w):
        """Front index (0 = non-dominated) of every row of ``w`` (N, M) weighted fitness values,
        larger is better; equal rows share a front (``gmm_nondominated_fronts``).  numpy in, numpy out."""
        w = np.ascontiguousarray(w, np.float64)
        n, m = w.shape
        if n == 0:
            return np.zeros(0, np.int64)
        order = np.lexsort(tuple(-w[:, c] for c in range(m - 1, -1, -1)))     # descending, column 0 first
        ws = torch.as_tensor(w[order]).to(self.torch_device)
        front = torch.empty(n, dtype=torch.int32, device=self.torch_device)
        _lib.check(self._L.gmm_nondominated_fronts(self._h, int(n), int(m), _dev_ptr(ws), _dev_ptr(front),
                                                   self._stream()))
        out = np.empty(n, np.int64)
        out[order] = front.cpu().numpy()
        return out

    def similar_lists(self, bits, n, row0=0, to_host=True):
        """Partner lists of a similarity bitmask from ``rmsd_matrix`` (rows ``row0 ...`` of the
        ``n x n`` matrix): ``(ptr int64 (rows + 1,), idx int32)``, self excluded; host arrays, or
        device tensors with ``to_host=False`` (what ``parallel.sharded_similar_lists`` gathers)."""
        rows = int(bits.shape[0])
        counts = torch.empty(rows, dtype=torch.int32, device=self.torch_device)
        _lib.check(self._L.gmm_similar_lists(self._h, int(n), int(row0), int(row0) + rows, _dev_ptr(bits),
                                             _dev_ptr(counts), None, None, 0, self._stream()))
        ptr = torch.zeros(rows + 1, dtype=torch.int64, device=self.torch_device)
        torch.cumsum(counts, 0, out=ptr[1:])
        total = int(ptr[-1].item())
        idx = torch.empty(max(total, 1), dtype=torch.int32, device=self.torch_device)
        _lib.check(self._L.gmm_similar_lists(self._h, int(n), int(row0), int(row0) + rows, _dev_ptr(bits),
                                             None, _dev_ptr(ptr), _dev_ptr(idx), total, self._stream()))
        if not to_host:
            return ptr, idx[:total]
        return ptr.cpu().numpy(), idx[:total].cpu().numpy()

    # -- geometric restraint objectives (Distance / Angle) -----------------------------------
    def set_geometry(self, terms, n_obj):
        """``terms``: dicts with ``kind`` (0 distance, 1 angle, 2 dihedral, 3 mass-weighted centre: point 0 a
        probe atom, ``threshold`` its mass, point 1 the target), ``obj`` (output column),
        ``points``: up to 4 entries, each a mobile atom index (int) or world coordinates (3 floats),
        ``threshold`` (float or ``'planar'`` for angles), optional ``tolerance``, ``weight``."""
        n = len(terms)
        kind = np.zeros(n, np.int32); obj = np.zeros(n, np.int32)
        atoms = -np.ones((n, 4), np.int32); points = np.zeros((n, 4, 3), np.float64)
        thr = np.zeros(n, np.float64); tol = np.full(n, np.nan); w = np.ones(n, np.float64)
        planar = np.zeros(n, np.int32)
        for k, t in enumerate(terms):
            kind[k], obj[k] = t["kind"], t["obj"]
            for j, pt in enumerate(t["points"]):
                if isinstance(pt, (int, np.integer)):
                    atoms[k, j] = int(pt)
                else:
                    points[k, j] = np.asarray(pt, np.float64)
            if t["threshold"] == "planar":
                planar[k] = 1
            else:
                thr[k] = float(t["threshold"])
            if t.get("tolerance") is not None:
                tol[k] = float(t["tolerance"])
            w[k] = float(t.get("weight", 1.0) if t.get("weight") is not None else 1.0)
        _lib.check(self._L.gmm_set_geometry(self._h, n, int(n_obj), _np_ptr(kind), _np_ptr(obj), _np_ptr(atoms),
                                            _np_ptr(points), _np_ptr(thr), _np_ptr(tol), _np_ptr(w), _np_ptr(planar)))
        self.n_geo_obj = int(n_obj)

    def geometry(self, xyz):
        """(P, n_obj) float64 restraint scores of expressed coordinates (P, M, 3)."""
        xyz = self._dev(xyz, torch.float32)
        P = int(xyz.shape[0])
        out = torch.empty((P, self.n_geo_obj), dtype=torch.float64, device=self.torch_device)
        _lib.check(self._L.gmm_geometry(self._h, P, _dev_ptr(xyz), _dev_ptr(out), self._stream()))
        return out

    def set_coordination(self, sites):
        """``sites``: one dict per Coordination objective with the whole-job atom tables
        (``mobile_row`` int32 [n], ``xyz`` float64 [n, 3], ``candidate`` / ``geometry`` uint8 [n],
        ``bond_radius`` float64 [n], ``neighbours``: list of bonded-atom lists in bond order),
        ``metal``, ``template`` [(v + 1), 3], ``radius``, ``distance``, ``weight``, ``min_atoms`` and the
        booleans ``prevent_intruders``, ``enforce_all_residues``, ``center_of_mass_correction``,
        ``distance_correction`` (gaudi/objectives/coordination.py:130-145)."""
        import ctypes as C
        arr = (_lib.CoordSite * max(1, len(sites)))()
        keep = []
        for k, s in enumerate(sites):
            n = len(s["mobile_row"])
            ptr = np.zeros(n + 1, np.int32)
            ptr[1:] = np.cumsum([len(v) for v in s["neighbours"]])
            idx = np.array([a for v in s["neighbours"] for a in v], np.int32)
            tabs = [np.ascontiguousarray(s["mobile_row"], np.int32), np.ascontiguousarray(s["xyz"], np.float64),
                    np.ascontiguousarray(s["candidate"], np.uint8), np.ascontiguousarray(s["geometry"], np.uint8),
                    np.ascontiguousarray(s["bond_radius"], np.float64), ptr, idx if len(idx) else np.zeros(1, np.int32),
                    np.ascontiguousarray(s["template"], np.float64)]
            keep.append(tabs)
            a = arr[k]
            a.n_atoms = n
            (a.mobile_row, a.xyz, a.candidate, a.geometry, a.bond_radius, a.nb_ptr, a.nb_idx,
             a.template_xyz) = [_np_ptr(t) for t in tabs]
            a.metal, a.n_vertices = int(s["metal"]), len(tabs[7]) - 1
            a.radius, a.distance, a.weight = float(s["radius"]), float(s["distance"]), float(s["weight"])
            a.min_atoms = int(s["min_atoms"])
            a.flags = ((_lib.COORD_PREVENT_INTRUDERS if s.get("prevent_intruders", True) else 0) |
                       (_lib.COORD_ENFORCE_ALL_RESIDUES if s.get("enforce_all_residues") else 0) |
                       (_lib.COORD_COM_CORRECTION if s.get("center_of_mass_correction") else 0) |
                       (_lib.COORD_DISTANCE_CORRECTION if s.get("distance_correction") else 0))
        _lib.check(self._L.gmm_set_coordination(self._h, len(sites), C.cast(arr, C.c_void_p)))
        self.n_coord_sites = len(sites)

    def coordination(self, xyz):
        """(P, n_sites) float64 Coordination scores of expressed coordinates (P, M, 3)."""
        xyz = self._dev(xyz, torch.float32)
        P = int(xyz.shape[0])
        out = torch.empty((P, self.n_coord_sites), dtype=torch.float64, device=self.torch_device)
        _lib.check(self._L.gmm_coordination(self._h, P, _dev_ptr(xyz), _dev_ptr(out), self._stream()))
        return out

    # -- parity / measurement helpers --------------------------------------------------
    def contact_pairs(self, xyz_one, max_pairs=1 << 16):
        """Pairs behind the Contacts score of one individual: list of
        (atom_a, atom_b, kind, overlap, value); static atoms are numbered M + index."""
        xyz = self._dev(xyz_one, torch.float32).reshape(self.mobile.n_atoms, 3)
        pairs = np.empty((max_pairs, 2), np.int32)
        kind = np.empty(max_pairs, np.int32)
        ov = np.empty(max_pairs, np.float64)
        val = np.empty(max_pairs, np.float64)
        n = C.c_int(0)
        _lib.check(self._L.gmm_contacts_pairs(self._h, _dev_ptr(xyz), max_pairs, _np_ptr(pairs), _np_ptr(kind),
                                              _np_ptr(ov), _np_ptr(val), C.byref(n), self._stream()))
        k = min(n.value, max_pairs)
        return [(int(pairs[i, 0]), int(pairs[i, 1]), int(kind[i]), float(ov[i]), float(val[i])) for i in range(k)]

    def set_profiling(self, on=True):
        _lib.check(self._L.gmm_set_profiling(self._h, int(bool(on))))

    def kernel_times(self):
        """Durations in ms of the last express / sweep / finalize kernels (profiling on)."""
        ms = (C.c_float * 3)()
        _lib.check(self._L.gmm_kernel_times(self._h, ms))
        return {"express": ms[0], "sweep": ms[1], "finalize": ms[2]}

    def last_pair_tests(self):
        """Pair tests executed by the last Contacts call; the one-kernel path keeps the tally only
        while ``set_profiling(True)`` is in effect (``gmm_last_pair_tests``)."""
        n = C.c_ulonglong(0)
        _lib.check(self._L.gmm_last_pair_tests(self._h, C.byref(n)))
        return int(n.value)


def measure_fp32_peak(device=0):
    """Measured dependent-free FFMA throughput in TFLOP/s (FP32 roofline denominator)."""
    v = C.c_double(0.0)
    _lib.check(_lib.lib().gmm_measure_fp32_peak(int(device), C.byref(v)))
    return float(v.value)


def launch_count():
    return int(_lib.lib().gmm_launch_count())


def niching_lists(ptr, idx, rank):
    """Kept candidates (indices, ascending) of the greedy niching pass over partner lists
    (``gmm_niching_host``); ``rank`` = lexicographic fitness rank, larger is fitter."""
    ptr = np.ascontiguousarray(ptr, np.int64); idx = np.ascontiguousarray(idx, np.int32)
    rank = np.ascontiguousarray(rank, np.int64)
    n = len(rank)
    kept = np.zeros(n, np.uint8)
    _lib.check(_lib.lib().gmm_niching_host(n, _np_ptr(ptr), _np_ptr(idx), _np_ptr(rank), _np_ptr(kept)))
    return np.flatnonzero(kept)


def bits_to_bool(bits, n):
    """(rows, ceil(n/32)) int32/uint32 words -> (rows, n) bool numpy array."""
    w = np.ascontiguousarray(bits.detach().cpu().numpy() if isinstance(bits, torch.Tensor) else bits)
    w = w.view(np.uint32)
    b = np.unpackbits(w.view(np.uint8).reshape(w.shape[0], -1), axis=1, bitorder="little")
    return b[:, :n].astype(bool)
