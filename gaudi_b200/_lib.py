"""ctypes binding of libgaudimm_b200.so (the C ABI declared in include/gaudimm_b200.h).

There is no CPU fallback: if the shared library has not been built (``python -c "import
__graft_entry__ as g; g.build()"``) importing this module's ``lib()`` raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgaudimm_b200.so")

GMM_OK, GMM_EINVAL, GMM_ECUDA, GMM_ESTATE, GMM_ELIMIT = 0, -1, -2, -3, -4
OPT_ONLY_INTERNAL, OPT_NO_CULL, OPT_TILED, OPT_SEQUENTIAL_EXPRESS, OPT_RMSD_SIMT, OPT_RMSD_HMMA = 1, 2, 4, 8, 16, 32
MAX_MOBILE_ATOMS, MAX_TORSIONS = 1024, 1024

_vp, _i, _d = C.c_void_p, C.c_int, C.c_double


class CoordSite(C.Structure):
    """``gmm_coord_site`` of include/gaudimm_b200.h."""
    _fields_ = [("n_atoms", _i), ("mobile_row", _vp), ("xyz", _vp), ("candidate", _vp), ("geometry", _vp),
                ("bond_radius", _vp), ("nb_ptr", _vp), ("nb_idx", _vp), ("metal", _i), ("n_vertices", _i),
                ("template_xyz", _vp), ("radius", _d), ("distance", _d), ("weight", _d), ("min_atoms", _i),
                ("flags", _i)]


COORD_PREVENT_INTRUDERS, COORD_ENFORCE_ALL_RESIDUES, COORD_COM_CORRECTION, COORD_DISTANCE_CORRECTION = 1, 2, 4, 8

# name -> (restype, argtypes); must list every symbol of include/gaudimm_b200.h
SIGNATURES = {
    "gmm_create": (_i, [C.POINTER(_vp), _i]),
    "gmm_destroy": (_i, [_vp]),
    "gmm_last_error": (C.c_char_p, []),
    "gmm_version": (_i, []),
    "gmm_reserve": (_i, [_vp, _i]),
    "gmm_reserve_rmsd": (_i, [_vp, _i, _i]),
    "gmm_set_mobile": (_i, [_vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i]),
    "gmm_set_rotamers": (_i, [_vp, _i, _vp, _vp]),
    "gmm_set_receptor": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gmm_set_params": (_i, [_vp, _d, _d, _d, _d, _d, _d, _d, _i]),
    "gmm_express": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp]),
    "gmm_contacts": (_i, [_vp, _i, _vp, _vp, _vp, _vp]),
    "gmm_evaluate": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gmm_evaluate_host": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gmm_evaluate_static": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gmm_express_is_tree": (_i, [_vp]),
    "gmm_exchange_create": (_i, [_vp, _i, _i, _i, _vp]),
    "gmm_exchange_connect": (_i, [_vp, _vp]),
    "gmm_exchange_publish": (_i, [_vp, _vp, _vp, _i, _i, C.POINTER(_vp), C.POINTER(_vp), _vp]),
    "gmm_evaluate_publish": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, C.POINTER(_vp), C.POINTER(_vp), _vp]),
    "gmm_exchange_state": (_i, [_vp, C.POINTER(C.c_uint), C.POINTER(_i), C.POINTER(_vp), C.POINTER(_vp)]),
    "gmm_exchange_destroy": (_i, [_vp]),
    "gmm_set_geometry": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gmm_geometry": (_i, [_vp, _i, _vp, _vp, _vp]),
    "gmm_set_coordination": (_i, [_vp, _i, _vp]),
    "gmm_coordination": (_i, [_vp, _i, _vp, _vp, _vp]),
    "gmm_rmsd_matrix": (_i, [_vp, _i, _i, _vp, _i, _i, _vp, _vp, _d, _vp]),
    "gmm_similar_lists": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, C.c_longlong, _vp]),
    "gmm_niching_host": (_i, [_i, _vp, _vp, _vp, _vp]),
    "gmm_fronts2_host": (_i, [_i, _vp, _vp]),
    "gmm_nondominated_fronts": (_i, [_vp, _i, _i, _vp, _vp, _vp]),
    "gmm_contacts_pairs": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _vp, C.POINTER(_i), _vp]),
    "gmm_measure_fp32_peak": (_i, [_i, C.POINTER(_d)]),
    "gmm_last_pair_tests": (_i, [_vp, C.POINTER(C.c_ulonglong)]),
    "gmm_set_profiling": (_i, [_vp, _i]),
    "gmm_kernel_times": (_i, [_vp, C.POINTER(C.c_float)]),
    "gmm_launch_count": (C.c_ulonglong, []),
}

_lib = None


class GmmError(RuntimeError):
    def __init__(self, code, message):
        RuntimeError.__init__(self, "libgaudimm_b200 error %d: %s" % (code, message))
        self.code = code


def lib():
    """The loaded library (loads on first use; raises if it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "%s not found: the CUDA extension has not been built and gaudi_b200 has no "
                "CPU fallback. Run `python -c 'import __graft_entry__ as g; g.build()'` "
                "from the repository root." % LIB_PATH)
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(code):
    if code != GMM_OK:
        raise GmmError(code, lib().gmm_last_error().decode("utf-8", "replace"))
    return code
