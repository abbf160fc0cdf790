"""gmm_evaluate_static: the evaluate entry point for hosts that are not allowed to allocate
device memory (a GPU shared with a framework that owns the memory pool, a latency-bound loop
that cannot take cudaMalloc's device-wide synchronisation).  Its scratch is an array built into
the library, shared by every call, so this checks two calls in a row — with different
populations, so that data left behind by the first cannot pass for the second — against the
regular gmm_evaluate.  Collected last on purpose (file name), and self-contained so that it
also passes alone in a fresh process."""
import ctypes as C

import numpy as np
import pytest

from gaudi_b200 import model as M
from gaudi_b200 import synth

pytestmark = pytest.mark.gpu


def test_static_scratch_evaluate_matches_regular_evaluate(built_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from gaudi_b200 import _lib
    from gaudi_b200.engine import Engine

    lig = synth.make_ligand(40, 10, seed=1)
    dev = torch.device("cuda", 0)

    def free_bytes():
        torch.cuda.synchronize()
        return torch.cuda.mem_get_info(dev)[0]

    first_call_on_device = True

    # Two Contacts paths use the built-in array differently: the default one-kernel path parks only
    # the expressed coordinates there, the tiled path (GMM_OPT_TILED) also its per-individual zone
    # bitmask and partial sums.  Both are checked, each on two calls in a row with different
    # populations, so that data left behind by the first cannot pass for the second.
    for params, n_rec, pops, too_big in ((M.ContactParams(), 3000, ((31, 1500), (32, 700)), 200000),
                                         (M.ContactParams(tiled=True), 5000, ((33, 4096), (34, 1000)), 20000)):
        rec = synth.make_receptor(n_rec, seed=2, params=params)
        eng = Engine(lig, rec, params, device=0)
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        for seed, pop in pops:
            tors, xf = synth.make_alleles(pop, lig.n_allele_torsions, seed=seed)
            want_s, want_c, _ = eng.evaluate(tors, xf)
            # inputs and outputs are the caller's: allocate them first, then watch the driver's
            # free-memory figure across the call itself (torch's caching allocator does not
            # return memory to the driver in between, so any drop would be the library's)
            t = torch.as_tensor(tors).to(dev)
            x = torch.as_tensor(xf).to(dev)
            scores = torch.full((pop, 2), float("nan"), dtype=torch.float64, device=dev)
            counts = torch.full((pop, 2), -1, dtype=torch.int32, device=dev)
            before = free_bytes()
            _lib.check(built_lib.gmm_evaluate_static(eng._h, pop, C.c_void_p(t.data_ptr()), C.c_void_p(x.data_ptr()),
                                                     None, C.c_void_p(scores.data_ptr()), C.c_void_p(counts.data_ptr()),
                                                     stream))
            drop = before - free_bytes()
            if first_call_on_device:
                # the one-time set-up: with CUDA's lazy module loading the driver places the library's
                # built-in array on the GPU when it is first touched (its 16 MiB plus the module's
                # code, never more than a few MiB); this is the cost the header says every GPU
                # that uses the library carries, paid once
                assert 0 <= drop <= (16 << 20) + (8 << 20), drop
                first_call_on_device = False
            else:
                assert drop == 0, "gmm_evaluate_static changed the amount of free device memory by %d bytes" % drop
            got_s, got_c = scores.cpu().numpy(), counts.cpu().numpy()
            want_s = want_s.cpu().numpy()
            # same tolerance as tests/test_gpu_parity.py (the sums are integer, so in fact equal)
            assert np.all(np.abs(got_s - want_s) <= 1e-4 * np.maximum(1.0, np.abs(want_s)))
            assert np.array_equal(got_c, want_c.cpu().numpy())
            assert np.isfinite(got_s).all() and (np.abs(got_s).sum() > 0)

        # a population that does not fit the built-in 16 MiB is refused, loudly, without allocating
        t = torch.zeros((too_big, lig.n_allele_torsions), dtype=torch.float64, device=dev)
        x = torch.zeros((too_big, 12), dtype=torch.float64, device=dev)
        s = torch.zeros((too_big, 2), dtype=torch.float64, device=dev)
        before = free_bytes()
        rc = built_lib.gmm_evaluate_static(eng._h, too_big, C.c_void_p(t.data_ptr()), C.c_void_p(x.data_ptr()), None,
                                           C.c_void_p(s.data_ptr()), None, stream)
        assert rc == _lib.GMM_ELIMIT
        assert b"built-in buffer" in built_lib.gmm_last_error()
        assert free_bytes() == before
        eng.close()
