#!/usr/bin/env python
"""Benchmark of the GaudiMM hot path on B200: individuals scored / s for
express (Torsion + Search) + Contacts (clashes and hydrophobic), plus RMSD pairs / s.

    python bench.py --gpus N --steps K --warmup W              # the B200 path (this repo)
    python bench.py --impl reference --gpus N --steps K ...    # the reference's CPU path

One "step" = one batched evaluate of the whole population (BASELINE.json configs[2]:
65,536 poses of a 40-atom / 10-torsion ligand against a 20,000-atom receptor box; with N
GPUs the population is sharded by individual and the fitness vectors are all-gathered).
Prints ONE JSON line on rank 0.  See DESIGN.md section "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (population, ligand atoms, torsions, receptor atoms)
    "c3": (65536, 40, 10, 20000),   # BASELINE.json configs[2] -- the headline
    "c2": (4096, 40, 10, 5000),     # BASELINE.json configs[1]
}
RMSD_N, RMSD_A = 16384, 60          # BASELINE.json configs[3]
FLOPS_PER_PAIR_TEST = 10
FLOPS_PER_RMSD_PAIR = 9 * RMSD_A    # 3A FADD + 3A FFMA


def build_system(name):
    from gaudi_b200 import model as M, synth
    pop, L, T, R = WORKLOADS[name]
    params = M.ContactParams()
    lig = synth.make_ligand(L, T, seed=1)
    rec = synth.make_receptor(R, seed=2, params=params)
    tors, xf = synth.make_alleles(pop, T, seed=3)
    return params, lig, rec, tors, xf


# ------------------------------------------------------------------------------------
# CPU baseline: the oracle (a numpy/scipy port of the reference's CPU path) on host cores
# ------------------------------------------------------------------------------------
_W = {}


def _cpu_init(name):
    import oracle_adapter as OA
    params, lig, rec, tors, xf = build_system(name)
    _W.update(OA=OA, params=params, lig=lig, rec=rec, tors=tors, xf=xf)


def _cpu_eval(rng):
    OA, lig = _W["OA"], _W["lig"]
    out = []
    for p in range(rng[0], rng[1]):
        x = OA.express(lig, _W["tors"][p], _W["xf"][p])
        r = OA.contacts(lig, _W["rec"], _W["params"], x, with_slack=False)
        out.append((r.clash, r.hydro))
    return out


class CpuPool(object):
    def __init__(self, name, cores):
        import multiprocessing as mp
        self.cores = cores
        for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
            os.environ.setdefault(var, "1")      # one process per core, no nested BLAS threads
        # spawn: the parent may hold a CUDA context, which must not be forked
        self.pool = mp.get_context("spawn").Pool(cores, initializer=_cpu_init, initargs=(name,))
        self.pool.map(_cpu_eval, [(0, 1)] * cores)       # warm every worker

    def run(self, start, count):
        per = max(1, count // (self.cores * 4))
        chunks = [(a, min(a + per, start + count)) for a in range(start, start + count, per)]
        t0 = time.perf_counter()
        res = self.pool.map(_cpu_eval, chunks)
        dt = time.perf_counter() - t0
        return dt, [s for c in res for s in c]

    def close(self):
        self.pool.close()
        self.pool.join()


# ------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        threading.Thread.__init__(self, daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._halt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        while not self._halt.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, nm in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.001)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------
def run_rmsd(args):
    """BASELINE.json configs[3]: similarity bitmask of 16,384 conformers x 60 atoms, row blocks over
    the ranks; coordinates are all-gathered first (each rank starts with its own block, as after a
    sharded evaluate), every rank computes its rows, the bitmask rows are all-gathered to rank 0."""
    import torch
    import torch.distributed as dist
    from gaudi_b200 import model as M, parallel, synth
    from gaudi_b200.engine import Engine, measure_fp32_peak, launch_count
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    lig = synth.make_ligand(12, 2, seed=1)
    eng = Engine(lig, M.empty_receptor(), M.ContactParams(only_internal=True), device=local)
    conf = synth.make_conformers(RMSD_N, RMSD_A, seed=4).astype(np.float32)
    b = parallel.shard_bounds(RMSD_N, world)
    mine = torch.as_tensor(conf[b[rank]:b[rank + 1]]).to(dev)
    flush = torch.empty(192 * 1024 * 1024, dtype=torch.uint8, device=dev)

    def rows(allc, r0, r1):
        return eng.rmsd_matrix(allc, row0=r0, row1=r1, threshold=2.5, values=False)[1]

    def step():
        return parallel.sharded_rmsd_bits(rows, mine, counts=[b[r + 1] - b[r] for r in range(world)])

    for _ in range(max(3, args.warmup)):
        bits = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    peak32 = measure_fp32_peak(local)
    l0 = launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for k in range(args.steps):
        flush.fill_(k & 0xff)
        ev[k][0].record(); step(); ev[k][1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t = torch.tensor([sum(a.elapsed_time(c) for a, c in ev) / args.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    if rank == 0:
        print(json.dumps({
            "metric": "RMSD pairs/s", "value": (RMSD_N * (RMSD_N - 1) // 2) / (ms * 1e-3), "unit": "unique pairs/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[3]: similarity bitmask of %d conformers x %d atoms, threshold 2.5 A, "
                                   "row-block sharded; fp16 tensor-core pre-filter, every pair inside its error bound "
                                   "re-decided by the exact f32 sum (bits identical to the f32 kernel)" % (RMSD_N, RMSD_A),
                       "l2": "192 MiB buffer written between timed iterations"},
            "roofline": dict(rmsd_roofline(((RMSD_N + 127) // 128) * (((RMSD_N + 127) // 128) + 1) // 2 * 128 * 128, ms, peak32),
                             step_note="step time includes the coordinate and bitmask all-gathers"),
            "gpu_launches": int(launch_count() - l0)}), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


def run_b200(args):
    import torch
    import torch.distributed as dist
    from gaudi_b200 import model as M
    from gaudi_b200.engine import Engine, measure_fp32_peak, launch_count
    from gaudi_b200 import synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    name = args.workload
    pop, L, T, R = WORKLOADS[name]
    params, lig, rec, tors, xf = build_system(name)
    # shard by individual: contiguous blocks, the last ranks may hold one fewer
    bounds = [(pop * r) // world for r in range(world + 1)]
    lo, hi = bounds[rank], bounds[rank + 1]
    shard = hi - lo
    max_shard = max(bounds[r + 1] - bounds[r] for r in range(world))

    eng = Engine(lig, rec, params, device=local)
    eng.reserve(max_shard)
    eng.set_profiling(True)
    t_d = torch.as_tensor(tors[lo:hi]).to(dev)
    x_d = torch.as_tensor(xf[lo:hi]).to(dev)
    scores = torch.empty((max_shard, 2), dtype=torch.float64, device=dev)
    counts = torch.empty((max_shard, 2), dtype=torch.int32, device=dev)
    gathered = torch.empty((world * max_shard, 2), dtype=torch.float64, device=dev) if world > 1 else None
    flush = torch.empty(192 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2
    # page-locked host buffers of the end-to-end leg (allocated up front, as an application would)
    th = torch.as_tensor(np.ascontiguousarray(tors[lo:hi])).pin_memory()
    xh = torch.as_tensor(np.ascontiguousarray(xf[lo:hi])).pin_memory()
    s_pin = torch.empty((shard, 2), dtype=torch.float64).pin_memory()
    c_pin = torch.empty((shard, 2), dtype=torch.int32).pin_memory()

    def step():
        eng.evaluate(t_d, x_d, counts=True, scores_out=scores[:shard], counts_out=counts[:shard])
        if world > 1:
            dist.all_gather_into_tensor(gathered, scores)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    peak32 = measure_fp32_peak(local)          # FP32 roofline denominator, measured here

    # Single GPU: the step (memset + express + Contacts) is captured once into a CUDA graph and
    # replayed (about 1 % at this size: the step is not launch-bound).  Scratch is reserved and the
    # kernels have run, so the captured calls neither allocate nor synchronise.  With several
    # ranks the step is launched eagerly: the NCCL all_gather follows the kernels on the same
    # stream and its enqueue is hidden behind them; capturing the collective bought nothing and
    # a two-rank trial of it did not complete.  --no-graph launches eagerly at N=1 too.
    graph, launches_per_step = None, None
    if not args.no_graph and world == 1:
        eng.set_profiling(False)               # event timing is for the eager passes below
        try:
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                step()
            torch.cuda.current_stream().wait_stream(side)
            barrier()
            graph = torch.cuda.CUDAGraph()
            n0 = launch_count()
            with torch.cuda.graph(graph):
                step()
            launches_per_step = launch_count() - n0
            graph.replay()
            barrier()
        except Exception as exc:               # capture unsupported (e.g. by the NCCL build): run eagerly
            sys.stderr.write("bench: CUDA graph capture failed (%s); launching eagerly\n" % exc)
            graph = None
            barrier()
    run_step = graph.replay if graph is not None else step
    eng.set_profiling(False)                   # no instrumentation inside the timed region
    for _ in range(3):
        run_step()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sweep_ms, expr_ms, fin_ms = [], [], []
    barrier()
    for k in range(args.steps):
        flush.fill_(k & 0xff)                  # evict L2 between timed iterations
        ev[k][0].record()
        run_step()
        ev[k][1].record()
    barrier()
    launches = launches_per_step * args.steps if graph is not None else launch_count() - launches0
    clocks = sampler.stop()
    # per-kernel durations (roofline numerator) and the executed pair tests from eager, profiled
    # launches of the same step, each kernel bracketed by CUDA events on the launching stream, L2
    # flushed as in the timed loop
    eng.set_profiling(True)
    for k in range(min(args.steps, 10)):
        flush.fill_(k & 0xff)
        step()
        kt = eng.kernel_times()
        sweep_ms.append(kt["sweep"]); expr_ms.append(kt["express"]); fin_ms.append(kt["finalize"])
    barrier()
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = pop / (ms_per_step * 1e-3)

    # roofline of the dominant kernel (contacts sweep), executed pair tests
    tests = eng.last_pair_tests()
    sweep = float(np.mean(sweep_ms))
    achieved = tests * FLOPS_PER_PAIR_TEST / (sweep * 1e-3) / 1e12
    dense_flops_per_ind = FLOPS_PER_PAIR_TEST * (L * R + L * (L - 1) // 2)

    # e2e: host buffers through the C ABI (H2D alleles, kernels, D2H scores) per step
    eng.set_profiling(False)
    for _ in range(max(3, args.warmup)):       # same warm-up as the device loop (also wakes the PCIe link)
        eng.evaluate_host(th, xh, scores_out=s_pin, counts_out=c_pin)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_e2e = max(3, args.steps)
    e0.record()
    for _ in range(n_e2e):
        s_h, c_h, _ = eng.evaluate_host(th, xh, scores_out=s_pin, counts_out=c_pin)
        if world > 1:
            dist.all_gather_into_tensor(gathered, scores)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1) / n_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item())
    h2d = int(th.numel() * th.element_size() + xh.numel() * xh.element_size())
    d2h = int(s_h.numel() * s_h.element_size() + c_h.numel() * c_h.element_size())
    assert torch.equal(s_pin.to(dev), scores[:shard]), "e2e path disagrees with the device path"

    extra = {}
    if rank == 0 and world == 1 and not args.quick:
        extra = side_measurements(eng, torch, dev, peak32)

    out = None
    if rank == 0:
        out = {
            "metric": "individuals scored/sec (expr + Contacts)", "value": value, "unit": "individuals/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "configs[2]: pop %d x (%d-atom, %d-torsion ligand) vs %d-atom receptor box, "
                                   "Torsion+Search express + Contacts(clashes, hydrophobic)" % (pop, L, T, R),
                       "sharding": "by individual, %d per GPU, NCCL all_gather of fitness" % max_shard if world > 1
                       else "single GPU",
                       "l2": "192 MiB buffer written between timed iterations",
                       "launch": "step replayed from a CUDA graph" if graph is not None else "eager launches",
                       "culling": "receptor tiles out of reach of a pose are skipped (results bit-identical to dense)"},
            "roofline": {"bound": "fp32", "kernel": "contacts_fused_kernel", "achieved": achieved, "peak": peak32,
                         "unit": "TFLOP/s", "frac": achieved / peak32 if peak32 else None,
                         # dram__bytes_read.sum + dram__bytes_write.sum of this kernel on this workload at
                         # N=1, from the committed capture profiles/ncu_contacts_fused_r01j.txt (bytes/launch)
                         "traffic": 31.76e6 if (name == "c3" and world == 1) else None,
                         "algorithmic_bytes_per_launch": shard * (L * 12 + 16 + 8),
                         "hbm": hbm_check(shard * (L * 12 + 16 + 8), sweep),
                         "note": "culled run: the pair-test flop count credits only the sweep (about a third of "
                                 "the issued instructions); culling, candidate expansion, pair scoring and the "
                                 "static-pair walk are per-individual bookkeeping. The dense run of the same "
                                 "kernel (key 'dense') is the sweep-dominated regime the FP32 roofline models. "
                                 "ncu of this kernel: profiles/ncu_contacts_fused_culled_r01.txt (issue slots "
                                 "~79 % busy, issue-bound)",
                         "peak_source": "measured here: dependent-free FFMA loop (MEASURED_PEAKS.json has no FP32 figure)",
                         "pair_tests_per_launch": tests, "flops_per_pair_test": FLOPS_PER_PAIR_TEST,
                         "kernel_ms": sweep, "express_ms": float(np.mean(expr_ms)), "finalize_ms": float(np.mean(fin_ms)),
                         "dense_flops_per_individual": dense_flops_per_ind,
                         "executed_fraction_of_dense": tests * FLOPS_PER_PAIR_TEST / (dense_flops_per_ind * shard)},
            "e2e": {"value": pop / (e2e_ms * 1e-3), "unit": "individuals/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms},
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
        out.update(extra)
    if rank == 0 and world == 1 and not args.no_cpu:
        out["cpu_baseline"] = cpu_baseline(name, args.cpu_sample)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out), flush=True)


def hbm_check(bytes_per_launch, kernel_ms):
    """HBM is not the bound of this path; say by how much, against the driver-measured copy bandwidth."""
    peak = None
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            peak = float(json.load(fh)["hbm_gbs"])
    except Exception:
        peak = 6650.0          # fallback stated in B200_PROFILING.md
    ach = bytes_per_launch / (kernel_ms * 1e-3) / 1e9
    return {"achieved_gbs": ach, "peak_gbs": peak, "frac": ach / peak}


def tensor_peak():
    """Dense fp16/bf16 tensor-core throughput to hold the HMMA work of the bitmask kernel against:
    the driver-measured cuBLAS bf16 burst figure, else the nominal one of B200_PROFILING.md."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["bf16_tflops"]), "MEASURED_PEAKS.json bf16_tflops (cuBLAS burst)"
    except Exception:
        return 2250.0, "nominal dense fp16/bf16 of B200_PROFILING.md (MEASURED_PEAKS.json absent)"


def rmsd_roofline(n_pairs_computed, ms, peak32):
    """The bitmask kernel does its arithmetic on the tensor cores (2 * 192 flops of HMMA per pair for
    60 atoms, K padded to 192) and is bounded by what follows the MMA: the per-pair decision, the
    exact re-decision of the pairs inside the error bound, bit packing.  Reported: the HMMA rate
    against the tensor peak, and the same pairs/s expressed in the FP32 kernel's flops (what the
    reference's formula costs) against the FP32 peak."""
    kp = (3 * RMSD_A + 15) // 16 * 16
    peak_t, src = tensor_peak()
    ach_t = n_pairs_computed * 2 * kp / (ms * 1e-3) / 1e12
    ach_eq = n_pairs_computed * FLOPS_PER_RMSD_PAIR / (ms * 1e-3) / 1e12
    return {"bound": "tensor", "kernel": "rmsd_bits_mma_kernel", "achieved": ach_t, "peak": peak_t, "unit": "TFLOP/s",
            "frac": ach_t / peak_t, "peak_source": src, "flops_per_pair": 2 * kp, "traffic": None,
            "fp32_equivalent": {"achieved": ach_eq, "peak": peak32, "frac": ach_eq / peak32 if peak32 else None,
                                "flops_per_pair": FLOPS_PER_RMSD_PAIR,
                                "note": "the FP32 direct-difference kernel (values requested, or GMM_OPT_RMSD_SIMT) "
                                        "reaches about 0.5 of the FP32 peak on the same pairs"},
            "note": "tensor pipe mostly idle by design: the MMA is a pre-filter, the kernel is bounded by the "
                    "decision epilogue (profiles/ncu_rmsd_bits_mma_r01.txt)"}


def side_measurements(eng, torch, dev, peak32):
    """Dense (no culling) figure on the same workload, configs[1], and the RMSD matrix of
    configs[3]; reported next to the headline, same timing rules."""
    from gaudi_b200 import model as M, synth
    from gaudi_b200.engine import Engine
    res = {}

    def timed(fn, reps):
        fn(); fn(); fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    # dense on the headline workload
    pop, L, T, R = WORKLOADS["c3"]
    dense = Engine(eng.mobile, eng.receptor, M.ContactParams(cull=False), device=eng.device)
    dense.set_profiling(True)
    tors, xf = synth.make_alleles(pop, T, seed=3)
    t_d, x_d = torch.as_tensor(tors).to(dev), torch.as_tensor(xf).to(dev)
    ms = timed(lambda: dense.evaluate(t_d, x_d), 3)
    kt = dense.kernel_times()
    tests = dense.last_pair_tests()
    ach = tests * FLOPS_PER_PAIR_TEST / (kt["sweep"] * 1e-3) / 1e12
    res["dense"] = {"value": pop / (ms * 1e-3), "unit": "individuals/s", "ms_per_step": ms,
                    "roofline": {"achieved": ach, "peak": peak32, "frac": ach / peak32, "unit": "TFLOP/s",
                                 "kernel_ms": kt["sweep"], "pair_tests_per_launch": tests}}
    dense.close()
    # configs[1]
    params, lig, rec, tors, xf = build_system("c2")
    e2 = Engine(lig, rec, params, device=eng.device)
    t_d, x_d = torch.as_tensor(tors).to(dev), torch.as_tensor(xf).to(dev)
    ms = timed(lambda: e2.evaluate(t_d, x_d), 20)
    res["configs1_pop4096_rec5k"] = {"value": tors.shape[0] / (ms * 1e-3), "unit": "individuals/s", "ms_per_step": ms}
    e2.close()
    # configs[3]: RMSD matrix, bitmask output (what niching consumes)
    conf = torch.as_tensor(synth.make_conformers(RMSD_N, RMSD_A, seed=4), dtype=torch.float32).to(dev)
    bits = torch.empty((RMSD_N, RMSD_N // 32), dtype=torch.int32, device=dev)
    from gaudi_b200 import _lib
    import ctypes as C

    def rm():
        _lib.check(eng._L.gmm_rmsd_matrix(eng._h, RMSD_N, RMSD_A, C.c_void_p(conf.data_ptr()), 0, RMSD_N, None,
                                          C.c_void_p(bits.data_ptr()), 2.5, eng._stream()))
    ms = timed(rm, 5)
    nt = (RMSD_N + 127) // 128
    ordered = nt * (nt + 1) // 2 * 128 * 128        # pairs the kernel computes (upper-triangle tiles)
    res["rmsd"] = {"metric": "RMSD pairs/s", "value": (RMSD_N * (RMSD_N - 1) // 2) / (ms * 1e-3),
                   "unit": "unique pairs/s", "ms": ms, "n_conformers": RMSD_N, "atoms": RMSD_A,
                   "note": "N x N similarity bitmask; tiles above the diagonal are computed once and mirrored; "
                           "value counts unique pairs, the roofline counts the pairs actually computed",
                   "roofline": rmsd_roofline(ordered, ms, peak32)}
    # the FP32 direct-difference kernel on the same conformers (what a call that wants the values runs)
    e_simt = Engine(eng.mobile, M.empty_receptor(), M.ContactParams(only_internal=True, rmsd_simt=True), device=dev.index or 0)

    def rm_simt():
        _lib.check(e_simt._L.gmm_rmsd_matrix(e_simt._h, RMSD_N, RMSD_A, C.c_void_p(conf.data_ptr()), 0, RMSD_N, None,
                                             C.c_void_p(bits.data_ptr()), 2.5, e_simt._stream()))
    ms_s = timed(rm_simt, 5)
    ach_s = ordered * FLOPS_PER_RMSD_PAIR / (ms_s * 1e-3) / 1e12
    res["rmsd"]["fp32_kernel"] = {"value": (RMSD_N * (RMSD_N - 1) // 2) / (ms_s * 1e-3), "unit": "unique pairs/s", "ms": ms_s,
                                  "roofline": {"bound": "fp32", "kernel": "rmsd_tile_kernel", "achieved": ach_s,
                                               "peak": peak32, "frac": ach_s / peak32, "unit": "TFLOP/s"}}
    e_simt.close()
    try:
        res["configs4_rotamers12_pop32768"] = rotamers_measurement(torch, dev, timed)
    except Exception as exc:                       # a side measurement must never sink the headline
        res["configs4_rotamers12_pop32768"] = {"error": repr(exc)[:200]}
    return res


def rotamers_measurement(torch, dev, timed):
    """BASELINE.json configs[4] on one GPU: Rotamers gene on 12 flexible side chains + Torsion +
    Search + Contacts, population 32,768, through the plugin layer's model builder.  Structure:
    the reference's own 5er1 test complex (committed fixture); chi tables: seeded random rows
    (the reference takes them from Chimera's Dunbrack library, which is not available here)."""
    import random
    import golden_util as G
    from gaudi_b200 import base
    cfg, _ = G.configs4_job(12, seed=5)
    lig = cfg.genes[1]["mol"]
    random.seed(5)
    template = base.MolecularIndividual(cfg)
    env = base.Environment(cfg, device=dev.index or 0)
    batch = env.attach(template)
    eng = batch.engines[0][0]
    pop = 32768
    tors, xf, rot = G.configs4_alleles(batch, template, pop, seed=5)
    t_d, x_d, r_d = (torch.as_tensor(v).to(dev) for v in (tors, xf, rot))
    eng.set_profiling(True)
    ms = timed(lambda: eng.evaluate(t_d, x_d, r_d), 10)
    kt = eng.kernel_times()
    out = {"value": pop / (ms * 1e-3), "unit": "individuals/s", "ms_per_step": ms, "population": pop,
           "express_ms": kt["express"], "contacts_ms": kt["sweep"],
           "mobile_atoms": eng.mobile.n_atoms, "ligand_atoms": lig.n_atoms, "torsions": batch.n_torsions,
           "rotamer_residues": batch.n_rot_res, "receptor_atoms": eng.receptor.n_atoms,
           "structure": "5er1 test complex (86-atom ligand, 4,669-atom all-H protein), chi tables synthetic"}
    batch.close()
    return out


def c_oracle_for(name):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import c_oracle
    params, lig, rec, tors, xf = build_system(name)
    return c_oracle.COracle(lig, rec, params), tors, xf


def cpu_baseline(name, sample):
    """The oracle's C restatement (fp64, cell list, every host core) on a bounded sample of the
    same population; the numpy restatement it is validated against is timed beside it."""
    co, tors, xf = c_oracle_for(name)
    sample = min(sample, len(tors))
    co.evaluate(tors[:256], xf[:256])                      # warm
    reps, dt = 0, 0.0
    while dt < 10.0 and reps < 40:                         # about 10 s of host work
        t0 = time.perf_counter()
        co.evaluate(tors[:sample], xf[:sample])
        dt += time.perf_counter() - t0
        reps += 1
    sample_total = sample * reps
    out = {"value": sample_total / dt, "unit": "individuals/s", "cores": co.workers, "kind": "port",
           "sample": "%d passes over the first %d individuals of the same population, C restatement of the reference's CPU path "
                     "(oracle/gaudi_oracle.c: fp64 express + zone + DetectClash rule + scoring, cell list), "
                     "%d host threads, %.1f s" % (reps, sample, co.workers, dt)}
    try:
        cores = os.cpu_count() or 1
        pool = CpuPool(name, cores)
        n_np = 64 * cores
        dt2, _ = pool.run(0, n_np)
        pool.close()
        out["numpy_port"] = {"value": n_np / dt2, "unit": "individuals/s", "cores": cores,
                             "sample": "%d individuals, numpy/scipy restatement (oracle/gaudi_oracle.py), %d processes" % (n_np, cores)}
    except Exception as exc:
        out["numpy_port"] = {"error": repr(exc)[:160]}
    return out


# ------------------------------------------------------------------------------------
# reference arm: the reference's CPU path (oracle port; Chimera itself cannot run here)
# ------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.workload
    pop, L, T, R = WORKLOADS[name]
    co, tors, xf = c_oracle_for(name)
    cores = co.workers
    # one step = one batch of the same population, sized from a probe so that a step is about
    # 1.5 s of host work (the whole 65,536-individual batch on a 16-core box)
    t0 = time.perf_counter()
    co.evaluate(tors[:4096], xf[:4096])
    speed = 4096 / (time.perf_counter() - t0)
    per_step = int(min(pop, max(cores * 64, speed * 1.5)))
    for w in range(min(args.warmup, 1)):
        co.evaluate(tors[:per_step], xf[:per_step])
    total = 0.0
    for k in range(args.steps):
        lo = (k * per_step) % max(1, pop - per_step + 1)
        t0 = time.perf_counter()
        co.evaluate(tors[lo:lo + per_step], xf[lo:lo + per_step])
        total += time.perf_counter() - t0
    value = per_step * args.steps / total
    sample = ("each step = %d individuals of the %d-individual population; C restatement (fp64, cell list) of "
              "the reference's Chimera-based CPU path (the reference itself needs UCSF Chimera + Python 2 and "
              "cannot run on this box), %d host threads" % (per_step, pop, cores))
    out = {"impl": "reference", "metric": "individuals scored/sec (expr + Contacts)", "value": value,
           "unit": "individuals/s", "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": total / args.steps * 1e3, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": "configs[2]: pop %d x (%d-atom, %d-torsion ligand) vs %d-atom receptor box, "
                                  "Torsion+Search express + Contacts(clashes, hydrophobic)" % (pop, L, T, R)},
           "cpu_baseline": {"value": value, "unit": "individuals/s", "cores": cores, "kind": "port", "sample": sample},
           "e2e": {"value": value, "unit": "individuals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS) + ["rmsd"])
    ap.add_argument("--cpu-sample", type=int, default=65536, help="individuals in the CPU baseline sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-graph", action="store_true", help="launch every step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--quick", action="store_true", help="skip the dense / configs[1] / RMSD side measurements")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.workload == "rmsd":
            args.workload = "c3"
        run_reference(args)
    elif args.workload == "rmsd":
        run_rmsd(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
