#!/usr/bin/env python
"""Benchmark of the GaudiMM hot path on B200: individuals scored / s for
express (Torsion + Search [+ Rotamers]) + Contacts (clashes and hydrophobic), plus RMSD pairs / s.

    python bench.py --gpus N --steps K --warmup W              # the B200 path (this repo)
    python bench.py --impl reference --gpus N --steps K ...    # the reference's CPU path
    (N > 1: python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...)

One "step" = one batched evaluate of the whole population (BASELINE.json configs[2]: 65,536
poses of a 40-atom / 10-torsion ligand against a 20,000-atom receptor box).  With N GPUs the
population is sharded by individual through the product's own multi-GPU layer
(gaudi_b200.parallel.ShardedEvaluator: every rank scores its block, the fitness rows are gathered
into every rank's table by stores through peer memory — no collective call in the step, the step
is one CUDA graph at every N; --exchange nccl selects an NCCL all_gather instead).  Prints ONE
JSON line on rank 0.  Side objects of the same line, at every N: "rmsd" (configs[3], row blocks
over the ranks, only the partner lists travel) and "configs4" (Rotamers on 12 side chains,
population 32,768, sharded by individual).  See DESIGN.md section 5 for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import re
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
RMSD_N, RMSD_A, RMSD_THR = 16384, 60, 2.5   # BASELINE.json configs[3]; threshold of gaudi/parse.py:279
C5_POP, C5_RES = 32768, 12          # BASELINE.json configs[4]
FLOPS_PER_PAIR_TEST = 10
FLOPS_PER_RMSD_PAIR = 9 * RMSD_A    # 3A FADD + 3A FFMA
L2_FLUSH_BYTES = 192 * 1024 * 1024  # > 126 MB L2

C3_TEXT = ("configs[2]: pop %d x (%d-atom, %d-torsion ligand) vs %d-atom receptor box, "
           "Torsion+Search express + Contacts(clashes, hydrophobic)" % WORKLOADS["c3"])
RMSD_TEXT = ("configs[3]: similarity (gaudi/similarity.py:36-75) of all pairs of %d conformers x %d atoms, threshold %.1f A"
             % (RMSD_N, RMSD_A, RMSD_THR))
C5_TEXT = ("configs[4]: Rotamers gene on %d flexible side chains + Torsion + Search + Contacts(clashes, hydrophobic), "
           "pop %d, 5er1 test complex" % (C5_RES, C5_POP))


def build_system(name):
    from gaudi_b200 import model as M, synth
    pop, L, T, R = WORKLOADS[name]
    params = M.ContactParams()
    lig = synth.make_ligand(L, T, seed=1)
    rec = synth.make_receptor(R, seed=2, params=params)
    tors, xf = synth.make_alleles(pop, T, seed=3)
    return params, lig, rec, tors, xf


def build_c5(device):
    """configs[4] through the plugin layer's model builder (tests/golden_util.configs4_job)."""
    import random
    import golden_util as G
    from gaudi_b200 import base
    cfg, _ = G.configs4_job(C5_RES, seed=5)
    random.seed(5)
    template = base.MolecularIndividual(cfg)
    env = base.Environment(cfg, device=device, group=False)
    batch = env.attach(template)
    tors, xf, rot = G.configs4_alleles(batch, template, C5_POP, seed=5)
    return batch, tors, xf, rot


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
# small helpers
# ------------------------------------------------------------------------------------
class Ctx(object):
    """Process-wide state of one bench run: ranks, device, L2 flush buffer, timing."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.args = torch, dist, args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)
        self.flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, values):
        t = self.torch.tensor(list(values), dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.cpu().numpy()

    def time_steps(self, fn, steps, warmup=3):
        """Per-step device times (ms): L2 flushed before every step, CUDA events on the launching
        stream, barrier + synchronize on both sides, per step the MAX over the ranks."""
        torch = self.torch
        for _ in range(warmup):
            fn()
        self.barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for k in range(steps):
            self.flush.fill_(k & 0xff)
            ev[k][0].record()
            fn()
            ev[k][1].record()
        self.barrier()
        return self.max_over_ranks(a.elapsed_time(b) for a, b in ev)


def spread(ms):
    return {"ms_per_step_min": float(np.min(ms)), "ms_per_step_median": float(np.median(ms)),
            "ms_per_step_max": float(np.max(ms))}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh)
    except Exception:
        return {}


def hbm_check(bytes_per_launch, kernel_ms):
    """HBM is not the bound of this path; say by how much, against the driver-measured copy bandwidth."""
    peak = float(measured_peaks().get("hbm_gbs", 6650.0))       # fallback stated in B200_PROFILING.md
    ach = bytes_per_launch / (kernel_ms * 1e-3) / 1e9
    return {"achieved_gbs": ach, "peak_gbs": peak, "frac": ach / peak}


def tensor_peak():
    pk = measured_peaks()
    if "bf16_tflops" in pk:
        return float(pk["bf16_tflops"]), "MEASURED_PEAKS.json bf16_tflops (cuBLAS burst)"
    return 2250.0, "nominal dense fp16/bf16 of B200_PROFILING.md (MEASURED_PEAKS.json absent)"


def committed_dram_traffic(kernel_tag):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of a kernel, read at run time from
    the newest committed ncu summary under profiles/ whose name holds ``kernel_tag``.  Returns
    (bytes or None, file name or None)."""
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    best = None
    pdir = os.path.join(ROOT, "profiles")
    for fn in sorted(os.listdir(pdir)) if os.path.isdir(pdir) else []:
        if kernel_tag in fn and fn.endswith(".txt"):
            best = fn
    if best is None:
        return None, None
    total, seen = 0.0, 0
    for line in open(os.path.join(pdir, best)):
        m = re.match(r"\s*dram__bytes_(read|write)\.sum\s+(\w+)\s+([0-9.,eE+-]+)", line)
        if m and m.group(2) in unit:
            total += float(m.group(3).replace(",", "")) * unit[m.group(2)]
            seen += 1
        if seen == 2:
            break
    return (total if seen == 2 else None), ("profiles/" + best)


# ------------------------------------------------------------------------------------
# CPU baselines (the oracle on the host cores; rank 0, N = 1 only)
# ------------------------------------------------------------------------------------
def c_oracle_module():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import c_oracle
    return c_oracle


def cpu_contacts_baseline(co, alleles, budget_s=10.0, what=""):
    """The oracle's C restatement (fp64, cell list, every host core) over the same population."""
    n = len(alleles[0])
    co.evaluate(*[a[:256] if a is not None else None for a in alleles])       # warm
    reps, dt = 0, 0.0
    while dt < budget_s and reps < 40:
        t0 = time.perf_counter()
        co.evaluate(*alleles)
        dt += time.perf_counter() - t0
        reps += 1
    return {"value": n * reps / dt, "unit": "individuals/s", "cores": co.workers, "kind": "port",
            "sample": "%d passes over %d individuals of the same population%s, C restatement of the reference's CPU path "
                      "(oracle/gaudi_oracle.c: fp64 express + zone + DetectClash rule + scoring, cell list), "
                      "%d host threads, %.1f s" % (reps, n, what, co.workers, dt)}


def cpu_rmsd_baseline(budget_s=8.0):
    """RMSD CPU baselines on a 1,024-conformer subsample of the same conformers (SURVEY.md 8d,
    BASELINE.md 3): (a) the reference's formula called pair by pair from Python, as
    gaudi/algorithms.py:141-155 calls ``similarity`` — the timed function is the oracle's
    ``rmsd_squared``, which restates gaudi/similarity.py:73-75 line for line and is pinned to
    recorded outputs of the reference's own module (tests/golden/similarity_reference.npz); the
    reference's file itself cannot be read on this box; (b) the vectorised numpy form of the same
    formula; (c) the C restatement on every host core."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import gaudi_oracle as O
    from gaudi_b200 import synth
    conf = synth.make_conformers(RMSD_N, RMSD_A, seed=4).astype(np.float32).astype(np.float64)
    sub = conf[:1024]
    thr2 = RMSD_THR * RMSD_THR
    # (a) pairwise, python loop
    t0, done, similar = time.perf_counter(), 0, 0
    for i in range(len(sub)):
        a = sub[i]
        for j in range(i + 1, len(sub)):
            similar += not (O.rmsd_squared(a, sub[j]) > thr2)
        done += len(sub) - 1 - i
        if time.perf_counter() - t0 > budget_s:
            break
    dt_a = time.perf_counter() - t0
    # (b) vectorised numpy, one row against the subsample at a time
    t0, done_b = time.perf_counter(), 0
    while time.perf_counter() - t0 < budget_s / 2:
        for i in range(0, len(sub), 8):
            d = sub - sub[i]
            _ = (d * d).sum(axis=(1, 2)) / RMSD_A > thr2
            done_b += len(sub)
    dt_b = time.perf_counter() - t0
    out = {"value": done / dt_a, "unit": "pairs/s", "cores": 1, "kind": "port",
           "sample": "%d pairs of a 1,024-conformer subsample of the same conformers, one Python call of the "
                     "reference's formula per pair as gaudi/algorithms.py:141-155 makes them (oracle.rmsd_squared = "
                     "gaudi/similarity.py:73-75, pinned to the reference module's recorded outputs), %.1f s"
                     % (done, dt_a),
           "numpy_vectorised": {"value": done_b / dt_b, "unit": "pairs/s", "cores": 1,
                                "sample": "same formula broadcast over one row of the subsample at a time, %.1f s" % dt_b}}
    try:
        co = c_oracle_module()
        lib = co._lib()
        import ctypes as C
        rows = 512
        buf = np.empty((rows, RMSD_N))
        nthreads = int(lib.go_num_threads())
        t0, reps = time.perf_counter(), 0
        while time.perf_counter() - t0 < budget_s / 2 or reps == 0:
            lib.go_rmsd_matrix(C.c_int(RMSD_N), C.c_int(RMSD_A), conf.ctypes.data_as(C.c_void_p), C.c_int(0),
                               C.c_int(rows), buf.ctypes.data_as(C.c_void_p))
            reps += 1
        dt_c = time.perf_counter() - t0
        out["c_port_all_cores"] = {"value": rows * RMSD_N * reps / dt_c, "unit": "pairs/s", "cores": nthreads,
                                   "sample": "%d passes over %d rows x %d columns, oracle/gaudi_oracle.c go_rmsd_matrix "
                                             "(fp64, OpenMP), %.1f s" % (reps, rows, RMSD_N, dt_c)}
    except Exception as exc:
        out["c_port_all_cores"] = {"error": repr(exc)[:160]}
    return out


# ------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------
def capture_graph(ctx, step):
    """The step as one CUDA graph (no allocation, no host synchronisation, no collective call
    inside).  Returns (graph or None, kernel launches per replay)."""
    torch = ctx.torch
    from gaudi_b200.engine import launch_count
    try:
        side = torch.cuda.Stream(device=ctx.dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            step()
        torch.cuda.current_stream().wait_stream(side)
        ctx.barrier()
        graph = torch.cuda.CUDAGraph()
        n0 = launch_count()
        with torch.cuda.graph(graph):
            step()
        per = launch_count() - n0
        graph.replay()
        ctx.barrier()
        return graph, per
    except Exception as exc:
        sys.stderr.write("bench: CUDA graph capture failed (%s); launching eagerly\n" % exc)
        ctx.barrier()
        return None, None


def bench_c3(ctx):
    torch, args = ctx.torch, ctx.args
    from gaudi_b200 import parallel
    from gaudi_b200.engine import Engine, measure_fp32_peak, launch_count
    name = "c3"
    pop, L, T, R = WORKLOADS[name]
    params, lig, rec, tors, xf = build_system(name)
    eng = Engine(lig, rec, params, device=ctx.local)
    sh = parallel.ShardedEvaluator(eng, pop, exchange=args.exchange)
    shard = sh.hi - sh.lo
    t_d = torch.as_tensor(sh.shard(tors)).to(ctx.dev)
    x_d = torch.as_tensor(sh.shard(xf)).to(ctx.dev)

    def step():
        return sh.evaluate(t_d, x_d)

    warm = max(3, args.warmup)
    for _ in range(warm):
        step()
    ctx.barrier()
    peak32 = measure_fp32_peak(ctx.local)          # FP32 roofline denominator, measured here
    graph, per_replay = (None, None)
    if not args.no_graph and sh.kind != "nccl":
        graph, per_replay = capture_graph(ctx, step)
    run_step = graph.replay if graph is not None else step
    sampler = ClockSampler(ctx.local)
    sampler.start()
    l0 = launch_count()
    ms = ctx.time_steps(run_step, args.steps, warmup=3)
    launches = per_replay * (args.steps + 3) if graph is not None else launch_count() - l0
    clocks = sampler.stop()
    ms_per_step = float(np.mean(ms))
    value = pop / (ms_per_step * 1e-3)

    # the gathered table every rank now holds equals the rows this rank computed itself
    if sh.exchange is not None:
        epoch, timed_out, g_s, g_c = sh.exchange.state()
        assert not timed_out, "fitness exchange: a wait for another rank timed out"
    else:
        g_s, g_c = step()
        torch.cuda.synchronize()
    own_s, own_c, _ = eng.evaluate(t_d, x_d)
    assert torch.equal(g_s[sh.lo:sh.hi], own_s) and torch.equal(g_c[sh.lo:sh.hi], own_c), "gathered rows differ"
    total = float(g_s[:pop].sum().item())
    checksum = ctx.max_over_ranks([total, -total])
    assert checksum[0] == -checksum[1], "ranks hold different fitness tables"

    # roofline of the dominant kernel: executed pair tests and kernel durations from eager, profiled
    # launches of the same step (CUDA events around each kernel on the launching stream, L2 flushed)
    eng.set_profiling(True)
    sweep_ms, expr_ms = [], []
    for k in range(min(args.steps, 10)):
        ctx.flush.fill_(k & 0xff)
        eng.evaluate(t_d, x_d, scores_out=sh.scores[:shard], counts_out=sh.counts[:shard])
        kt = eng.kernel_times()
        sweep_ms.append(kt["sweep"]); expr_ms.append(kt["express"])
    tests = eng.last_pair_tests()
    eng.set_profiling(False)
    ctx.barrier()
    sweep = float(np.mean(sweep_ms))
    achieved = tests * FLOPS_PER_PAIR_TEST / (sweep * 1e-3) / 1e12
    dense_flops_per_ind = FLOPS_PER_PAIR_TEST * (L * R + L * (L - 1) // 2)
    traffic, traffic_file = committed_dram_traffic("contacts_fused_culled") if (ctx.world == 1) else (None, None)

    e2e = bench_c3_e2e(ctx, eng, tors, xf, own_s, sh)

    out = None
    if ctx.rank == 0:
        out = {
            "metric": "individuals scored/sec (expr + Contacts)", "value": value, "unit": "individuals/s",
            "n_gpus": ctx.world, "steps": args.steps, "warmup": warm, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": C3_TEXT,
                       "sharding": ("by individual, %d per GPU; fitness rows gathered into every rank's table by %s"
                                    % (sh.max_shard, "stores through peer memory + epoch flags (gmm_exchange_publish), no collective call"
                                       if sh.kind == "peer" else "NCCL all_gather_into_tensor")) if ctx.world > 1 else "single GPU",
                       "l2": "192 MiB buffer written between timed iterations",
                       "launch": "step replayed from a CUDA graph" if graph is not None else "eager launches",
                       "culling": "receptor tiles out of reach of a pose are skipped (results bit-identical to dense)"},
            "roofline": {"bound": "fp32", "kernel": "contacts_fused_kernel", "achieved": achieved, "peak": peak32,
                         "unit": "TFLOP/s", "frac": achieved / peak32 if peak32 else None,
                         "traffic": traffic, "traffic_source": traffic_file,
                         "algorithmic_bytes_per_launch": shard * (L * 12 + 16 + 8),
                         "hbm": hbm_check(shard * (L * 12 + 16 + 8), sweep),
                         "note": "culled run: the pair-test flop count (SURVEY.md 8d: 10 flops per pair test) credits only "
                                 "the sweep; culling, candidate expansion, pair scoring and the static-pair walk are "
                                 "per-individual bookkeeping.  The dense run of the same kernel (key 'dense', N = 1) is "
                                 "the sweep-dominated regime the FP32 roofline models; the kernel executes 7 flops per "
                                 "pair test (3 FFMA + half a 3-input min), so by executed FP32 work the fractions are "
                                 "0.7 x these",
                         "peak_source": "measured here: dependent-free FFMA loop (MEASURED_PEAKS.json has no FP32 figure)",
                         "pair_tests_per_launch": tests, "flops_per_pair_test": FLOPS_PER_PAIR_TEST,
                         "kernel_ms": sweep, "express_ms": float(np.mean(expr_ms)),
                         "dense_flops_per_individual": dense_flops_per_ind,
                         "executed_fraction_of_dense": tests * FLOPS_PER_PAIR_TEST / (dense_flops_per_ind * shard)},
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
        out.update(spread(ms))
    state = {"eng": eng, "sh": sh, "peak32": peak32, "lig": lig, "rec": rec, "params": params, "tors": tors, "xf": xf}
    return out, state


def bench_c3_e2e(ctx, eng, tors, xf, own_scores, sh):
    """The same metric through the host-buffer API: the population's alleles start in page-locked
    HOST memory of rank 0 (the process that runs the GA), the fitness table ends there.  N = 1:
    gmm_evaluate_host.  N > 1: gaudi_b200.parallel.HostShardedEvaluator — the table is a
    shared-memory segment page-locked by every rank, each rank's gmm_evaluate_host uploads its rows
    over its own PCIe link and downloads its scores into the same table; rank 0 waits for every
    rank's done mark.  Timed with the host clock on rank 0 around the whole collective call
    (includes every copy and every wait), after a barrier."""
    torch, args = ctx.torch, ctx.args
    from gaudi_b200 import parallel
    pop = len(tors)
    hs = parallel.HostShardedEvaluator(eng, pop)
    if ctx.rank == 0:
        hs.tors[:pop] = tors
        hs.xforms[:pop] = xf
    n_e2e = max(5, args.steps)
    warm = max(3, args.warmup)
    times = []
    for k in range(warm + n_e2e):
        ctx.barrier()
        t0 = time.perf_counter()
        hs.evaluate(pop)
        dt = time.perf_counter() - t0
        if k >= warm:
            times.append(dt * 1e3)
    ctx.barrier()
    out = None
    if ctx.rank == 0:
        got = torch.as_tensor(hs.scores[sh.lo:sh.hi].copy()).to(ctx.dev)
        assert torch.equal(got, own_scores), "e2e path disagrees with the device path"
        assert np.isfinite(hs.scores[:pop]).all()
        e2e_ms = float(np.mean(times))
        row_in = hs.tors.shape[1] * 8 + 12 * 8
        out = {"value": pop / (e2e_ms * 1e-3), "unit": "individuals/s",
               "h2d_bytes_per_step": int(pop * row_in), "d2h_bytes_per_step": int(pop * (16 + 8)),
               "ms_per_step": e2e_ms, "ms_per_step_min": float(np.min(times)), "ms_per_step_max": float(np.max(times)),
               "timed_with": "host clock on rank 0 around the collective call, barrier before each step",
               "path": "gmm_evaluate_host" if ctx.world == 1 else
                       "parallel.HostShardedEvaluator: shared page-locked table, %d PCIe links in parallel" % ctx.world}
    hs.close()
    return out


def bench_rmsd(ctx, state, cpu=False):
    """configs[3]: every rank holds its block of conformers (as after a sharded evaluate); a step =
    all-gather of the conformers + this rank's rows of the similarity bitmask + compaction into
    partner lists + gather of the lists (what the niching pass reads)."""
    torch = ctx.torch
    import ctypes as C
    from gaudi_b200 import _lib
    from gaudi_b200 import model as M, parallel, synth
    from gaudi_b200.engine import Engine, launch_count
    lig = synth.make_ligand(12, 2, seed=1)
    eng = Engine(lig, M.empty_receptor(), M.ContactParams(only_internal=True), device=ctx.local)
    eng.reserve_rmsd(RMSD_N, RMSD_A)
    conf = synth.make_conformers(RMSD_N, RMSD_A, seed=4).astype(np.float32)
    b = parallel.shard_bounds(RMSD_N, ctx.world)
    counts = [b[r + 1] - b[r] for r in range(ctx.world)]
    mine = torch.as_tensor(conf[b[ctx.rank]:b[ctx.rank + 1]]).to(ctx.dev)
    allc = torch.as_tensor(conf).to(ctx.dev)
    res = {}
    sim = parallel.ShardedSimilarity(eng, counts, RMSD_A, threshold=RMSD_THR)

    def step():
        res["lists"] = sim.step(mine, root_only=True)      # the niching pass that reads them runs on rank 0

    l0 = launch_count()
    ms = ctx.time_steps(step, 10, warmup=3)
    launches = launch_count() - l0
    # the bitmask kernel alone, this rank's rows (roofline numerator)
    r0, r1 = b[ctx.rank], b[ctx.rank + 1]
    bits = torch.empty((r1 - r0, (RMSD_N + 31) // 32), dtype=torch.int32, device=ctx.dev)

    def kernel_only():
        _lib.check(eng._L.gmm_rmsd_matrix(eng._h, RMSD_N, RMSD_A, C.c_void_p(allc.data_ptr()), r0, r1, None,
                                          C.c_void_p(bits.data_ptr()), RMSD_THR, eng._stream()))
    kms = ctx.time_steps(kernel_only, 10, warmup=3)
    ptr, idx = res["lists"]
    n_pairs = RMSD_N * (RMSD_N - 1) // 2
    # pairs the kernels compute: 128 x 128 tiles; inside a rank's diagonal square only the upper triangle
    nt = (RMSD_N + 127) // 128
    rows_t = [((b[r + 1] - b[r]) + 127) // 128 for r in range(ctx.world)]
    computed = sum((rt * nt - rt * (rt - 1) // 2) * 128 * 128 for rt in rows_t)
    kp = (3 * RMSD_A + 15) // 16 * 16
    peak_t, src = tensor_peak()
    kernel_ms = float(np.mean(kms))
    ach_t = (computed / ctx.world) * 2 * kp / (kernel_ms * 1e-3) / 1e12
    out = None
    if ctx.rank == 0:
        step_ms = float(np.mean(ms))
        out = {"metric": "RMSD pairs/s", "value": n_pairs / (step_ms * 1e-3), "unit": "unique pairs/s",
               "n_gpus": ctx.world, "ms_per_step": step_ms, "scaling": "strong",
               "config": {"workload": RMSD_TEXT,
                          "step": ("all-gather of conformers (NCCL) + row-block bitmask + partner-list compaction + all-gather "
                                   "of the packed lists + copy of the counts, then of the entries, to rank 0's host memory "
                                   "(parallel.ShardedSimilarity, root_only)") if ctx.world > 1
                                  else "bitmask + partner-list compaction + copy of the lists to the host",
                          "l2": "192 MiB buffer written between timed iterations"},
               "partner_entries": int(idx.shape[0]),
               "partner_list_bytes": int(idx.shape[0] * 4 + ptr.shape[0] * 8),
               "bitmask_bytes": RMSD_N * ((RMSD_N + 31) // 32) * 4,
               "roofline": {"bound": "tensor", "kernel": "rmsd_bits kernel", "achieved": ach_t, "peak": peak_t,
                            "unit": "TFLOP/s", "frac": ach_t / peak_t, "peak_source": src, "flops_per_pair": 2 * kp,
                            "kernel_ms": kernel_ms, "pairs_computed_per_rank": computed // ctx.world,
                            "traffic": committed_dram_traffic("rmsd_bits_tc")[0] if ctx.world == 1 else None,
                            "traffic_source": committed_dram_traffic("rmsd_bits_tc")[1] if ctx.world == 1 else None,
                            "note": "MMA flops of the pairs this rank's tiles compute over the kernel's own duration "
                                    "(centring kernels included); the product is a pre-filter, pairs inside its error "
                                    "bound are re-decided by the exact f32 sum (bits identical to the FP32 kernel)"},
               "gpu_launches": int(launches)}
        out.update(spread(ms))
        if cpu:
            out["cpu_baseline"] = cpu_rmsd_baseline()
    # the niching pass of a population of 65,536 compares mu + lambda = 131,072 candidates: the size at
    # which row-block sharding pays (the 16,384 step above is bound by its fixed costs at any N)
    if not ctx.args.quick:
        try:
            big = rmsd_large(ctx, eng)
        except Exception as exc:
            big = {"error": repr(exc)[:300]}
        if ctx.rank == 0:
            out["n131072"] = big
    if ctx.world == 1 and ctx.rank == 0 and not ctx.args.quick:
        # the FP32 direct-difference kernel on the same conformers (what a call that wants the values runs)
        e_simt = Engine(lig, M.empty_receptor(), M.ContactParams(only_internal=True, rmsd_simt=True), device=ctx.local)

        def simt():
            _lib.check(e_simt._L.gmm_rmsd_matrix(e_simt._h, RMSD_N, RMSD_A, C.c_void_p(allc.data_ptr()), 0, RMSD_N, None,
                                                 C.c_void_p(bits.data_ptr()), RMSD_THR, e_simt._stream()))
        sms = float(np.mean(ctx.time_steps(simt, 5, warmup=2)))
        ach_s = computed * FLOPS_PER_RMSD_PAIR / (sms * 1e-3) / 1e12
        out["fp32_kernel"] = {"value": n_pairs / (sms * 1e-3), "unit": "unique pairs/s", "ms": sms,
                              "roofline": {"bound": "fp32", "kernel": "rmsd_tile_kernel", "achieved": ach_s,
                                           "peak": state["peak32"], "frac": ach_s / state["peak32"], "unit": "TFLOP/s",
                                           "flops_per_pair": FLOPS_PER_RMSD_PAIR}}
        e_simt.close()
    eng.close()
    return out


def rmsd_large(ctx, eng, n=131072):
    torch = ctx.torch
    from gaudi_b200 import parallel, synth
    b = parallel.shard_bounds(n, ctx.world)
    counts = [b[r + 1] - b[r] for r in range(ctx.world)]
    conf = synth.make_conformers(n, RMSD_A, seed=4)[b[ctx.rank]:b[ctx.rank + 1]].astype(np.float32)
    mine = torch.as_tensor(conf).to(ctx.dev)
    sim = parallel.ShardedSimilarity(eng, counts, RMSD_A, threshold=RMSD_THR, cap=64 * max(counts))
    res = {}

    def step():
        res["lists"] = sim.step(mine, root_only=True)
    ms = ctx.time_steps(step, 3, warmup=2)
    ptr, idx = res["lists"]
    step_ms = float(np.mean(ms))
    del sim
    torch.cuda.empty_cache()
    return {"n_conformers": n, "value": (n * (n - 1) // 2) / (step_ms * 1e-3), "unit": "unique pairs/s",
            "ms_per_step": step_ms, "partner_entries": int(len(idx)) if idx is not None else None, "n_gpus": ctx.world}


def bench_c5(ctx, state, cpu=False):
    """configs[4]: population 32,768 sharded by individual over the ranks, same exchange as the headline."""
    torch = ctx.torch
    from gaudi_b200 import parallel
    batch, tors, xf, rot = build_c5(ctx.local)
    eng = batch.engines[0][0]
    sh = parallel.ShardedEvaluator(eng, C5_POP, exchange=ctx.args.exchange)
    t_d, x_d, r_d = (torch.as_tensor(sh.shard(v)).to(ctx.dev) for v in (tors, xf, rot))

    def step():
        return sh.evaluate(t_d, x_d, r_d)

    for _ in range(3):
        step()
    ctx.barrier()
    graph, _ = (None, None) if (ctx.args.no_graph or sh.kind == "nccl") else capture_graph(ctx, step)
    ms = ctx.time_steps(graph.replay if graph is not None else step, 10, warmup=3)
    eng.set_profiling(True)
    km, em = [], []
    for k in range(5):
        ctx.flush.fill_(k)
        eng.evaluate(t_d, x_d, r_d)
        kt = eng.kernel_times()
        km.append(kt["sweep"]); em.append(kt["express"])
    tests = eng.last_pair_tests()
    eng.set_profiling(False)
    ctx.barrier()
    out = None
    if ctx.rank == 0:
        step_ms = float(np.mean(ms))
        kernel_ms = float(np.mean(km))
        ach = tests * FLOPS_PER_PAIR_TEST / (kernel_ms * 1e-3) / 1e12
        out = {"metric": "individuals scored/sec (expr + Contacts)", "value": C5_POP / (step_ms * 1e-3),
               "unit": "individuals/s", "n_gpus": ctx.world, "ms_per_step": step_ms, "scaling": "strong",
               "config": {"workload": C5_TEXT, "mobile_atoms": eng.mobile.n_atoms, "torsions": batch.n_torsions,
                          "rotamer_residues": batch.n_rot_res, "receptor_atoms": eng.receptor.n_atoms,
                          "structure": "5er1 test complex (86-atom ligand, all-H protein), chi tables synthetic",
                          "launch": "step replayed from a CUDA graph" if graph is not None else "eager launches",
                          "l2": "192 MiB buffer written between timed iterations"},
               "roofline": {"bound": "fp32", "kernel": "contacts_fused_kernel", "achieved": ach, "peak": state["peak32"],
                            "unit": "TFLOP/s", "frac": ach / state["peak32"], "kernel_ms": kernel_ms,
                            "express_ms": float(np.mean(em)), "pair_tests_per_launch": tests,
                            "flops_per_pair_test": FLOPS_PER_PAIR_TEST,
                            "traffic": committed_dram_traffic("contacts_fused_configs4")[0] if ctx.world == 1 else None,
                            "traffic_source": committed_dram_traffic("contacts_fused_configs4")[1] if ctx.world == 1 else None,
                            "note": "pair tests = mobile atoms x receptor atoms of the swept leaves + intra-mobile pairs"}}
        out.update(spread(ms))
        if cpu:
            co = c_oracle_module().COracle(eng.mobile, eng.receptor, eng.params)
            n = 8192
            out["cpu_baseline"] = cpu_contacts_baseline(co, (tors[:n], xf[:n], rot[:n]), budget_s=8.0,
                                                        what=" (first %d of %d)" % (n, C5_POP))
    sh.close()
    batch.close()
    return out


def bench_dense_and_c2(ctx, state):
    """N = 1 only: the dense (no culling) run of the headline workload and configs[1]."""
    torch = ctx.torch
    from gaudi_b200 import model as M
    from gaudi_b200.engine import Engine
    res = {}
    pop, L, T, R = WORKLOADS["c3"]
    dense = Engine(state["lig"], state["rec"], M.ContactParams(cull=False), device=ctx.local)
    dense.set_profiling(True)
    t_d, x_d = torch.as_tensor(state["tors"]).to(ctx.dev), torch.as_tensor(state["xf"]).to(ctx.dev)
    ms = float(np.mean(ctx.time_steps(lambda: dense.evaluate(t_d, x_d), 3, warmup=2)))
    kt = dense.kernel_times()
    tests = dense.last_pair_tests()
    ach = tests * FLOPS_PER_PAIR_TEST / (kt["sweep"] * 1e-3) / 1e12
    res["dense"] = {"value": pop / (ms * 1e-3), "unit": "individuals/s", "ms_per_step": ms,
                    "roofline": {"achieved": ach, "peak": state["peak32"], "frac": ach / state["peak32"], "unit": "TFLOP/s",
                                 "kernel_ms": kt["sweep"], "pair_tests_per_launch": tests,
                                 "note": "in SURVEY.md's 10-flop units; by executed FP32 work (7 flops per pair test) 0.7 x this"}}
    dense.close()
    params, lig, rec, tors, xf = build_system("c2")
    e2 = Engine(lig, rec, params, device=ctx.local)
    t_d, x_d = torch.as_tensor(tors).to(ctx.dev), torch.as_tensor(xf).to(ctx.dev)
    ms = float(np.mean(ctx.time_steps(lambda: e2.evaluate(t_d, x_d), 20, warmup=3)))
    res["configs1_pop4096_rec5k"] = {"value": tors.shape[0] / (ms * 1e-3), "unit": "individuals/s", "ms_per_step": ms}
    e2.close()
    return res


def bench_coordination(ctx):
    """N = 1 only: the Coordination objective on the reference's own test complex
    (tests/golden/4c3w_mol2.npz: vanadium site, octahedral template), 65,536 poses of the metal complex
    near the crystal pose; device-resident coordinates in, scores out.  CPU leg: the numpy oracle on a sample."""
    import time
    torch = ctx.torch
    import oracle_adapter as OA
    import test_gpu_coordination as TC
    mol, site, g, eng, mob = TC.complex_engine()
    pop = 65536
    xyz = eng.express(np.zeros((pop, 0)), TC.poses(mob, pop, seed=7))
    ms = float(np.mean(ctx.time_steps(lambda: eng.coordination(xyz), 10, warmup=3)))
    out = {"metric": "individuals scored/sec (Coordination objective)", "value": pop / (ms * 1e-3), "unit": "individuals/s",
           "ms_per_step": ms, "config": {"workload": "4c3w vanadium site, octahedron, radius 3.0, %d poses" % pop,
                                         "kernels": "coord_sphere_kernel + coord_cpd_kernel, fp64"}}
    x64 = xyz[:200].cpu().numpy().astype(np.float64) + eng.centre
    t0 = time.perf_counter()
    for p in range(200):
        OA.coordination(site, OA.coordination_world(site, x64[p]))
    out["cpu_baseline"] = {"value": 200 / (time.perf_counter() - t0), "unit": "individuals/s", "cores": 1, "kind": "port",
                           "sample": "200 poses, numpy restatement (oracle/coordination_oracle.py)"}
    eng.close()
    return out


def run_b200(args):
    ctx = Ctx(args)
    out, state = bench_c3(ctx)
    first = ctx.rank == 0 and ctx.world == 1
    if not args.quick:
        for key, fn in (("rmsd", bench_rmsd), ("configs4", bench_c5)):
            try:
                side = fn(ctx, state, cpu=first and not args.no_cpu)
            except Exception as exc:                   # a side measurement must never sink the headline
                side = {"error": repr(exc)[:300]}
            if ctx.rank == 0:
                out[key] = side
        if first:
            try:
                out.update(bench_dense_and_c2(ctx, state))
            except Exception as exc:
                out["dense"] = {"error": repr(exc)[:300]}
            if not args.no_cpu:
                try:
                    out["coordination"] = bench_coordination(ctx)
                except Exception as exc:
                    out["coordination"] = {"error": repr(exc)[:300]}
    if first and not args.no_cpu:
        co = c_oracle_module().COracle(state["lig"], state["rec"], state["params"])
        n = min(args.cpu_sample, len(state["tors"]))
        out["cpu_baseline"] = cpu_contacts_baseline(co, (state["tors"][:n], state["xf"][:n]))
    state["sh"].close()
    if ctx.world > 1:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out), flush=True)


def run_rmsd_only(args):
    ctx = Ctx(args)
    from gaudi_b200.engine import measure_fp32_peak
    state = {"peak32": measure_fp32_peak(ctx.local)}
    side = bench_rmsd(ctx, state, cpu=(ctx.world == 1 and not args.no_cpu))
    if ctx.rank == 0:
        line = {"metric": side["metric"], "value": side["value"], "unit": side["unit"], "n_gpus": ctx.world, "steps": 10,
                "warmup": 3, "ms_per_step": side["ms_per_step"], "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
        line.update({k: v for k, v in side.items() if k not in line})
        print(json.dumps(line), flush=True)
    if ctx.world > 1:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()


# ------------------------------------------------------------------------------------
# reference arm: the reference's CPU path (oracle port; Chimera itself cannot run here)
# ------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.workload == "rmsd":
        base = cpu_rmsd_baseline(budget_s=max(4.0, min(30.0, 2.0 * args.steps)))
        value = base["value"]
        out = {"impl": "reference", "metric": "RMSD pairs/s", "value": value, "unit": "pairs/s", "n_gpus": world,
               "steps": 1, "warmup": 0, "ms_per_step": None, "higher_is_better": True, "scaling": "strong",
               "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": {"workload": RMSD_TEXT},
               "cpu_baseline": base,
               "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
               "gpu_launches": 0}
        print(json.dumps(out), flush=True)
        return
    name = args.workload
    pop, L, T, R = WORKLOADS[name]
    params, lig, rec, tors, xf = build_system(name)
    co = c_oracle_module().COracle(lig, rec, params)
    cores = co.workers
    # one step = one batch of the same population, sized from a probe so that a step is about
    # 1.5 s of host work (the whole 65,536-individual batch on a 16-core box)
    t0 = time.perf_counter()
    co.evaluate(tors[:4096], xf[:4096])
    speed = 4096 / (time.perf_counter() - t0)
    per_step = int(min(pop, max(cores * 64, speed * 1.5)))
    warm = min(args.warmup, 2)                    # the host path has no clocks or caches to ramp: two passes at most
    for w in range(warm):
        co.evaluate(tors[:per_step], xf[:per_step])
    times = []
    for k in range(args.steps):
        lo = (k * per_step) % max(1, pop - per_step + 1)
        t0 = time.perf_counter()
        co.evaluate(tors[lo:lo + per_step], xf[lo:lo + per_step])
        times.append(time.perf_counter() - t0)
    total = sum(times)
    value = per_step * args.steps / total
    sample = ("each step = %d individuals of the %d-individual population; C restatement (fp64, cell list) of "
              "the reference's Chimera-based CPU path (the reference itself needs UCSF Chimera + Python 2 and "
              "cannot run on this box), %d host threads" % (per_step, pop, cores))
    out = {"impl": "reference", "metric": "individuals scored/sec (expr + Contacts)", "value": value,
           "unit": "individuals/s", "n_gpus": world, "steps": args.steps,
           "warmup": warm, "warmup_requested": args.warmup, "ms_per_step": total / args.steps * 1e3, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": C3_TEXT if name == "c3" else "configs[1]: pop %d, %d-atom ligand, %d-atom receptor" % (pop, L, R)},
           "cpu_baseline": {"value": value, "unit": "individuals/s", "cores": cores, "kind": "port", "sample": sample},
           "e2e": {"value": value, "unit": "individuals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    out.update(spread(np.array(times) * 1e3))
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c3", "c2", "rmsd"])
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1: how the fitness rows are gathered (peer-memory stores, or an NCCL all_gather)")
    ap.add_argument("--cpu-sample", type=int, default=65536, help="individuals in the CPU baseline sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline legs")
    ap.add_argument("--no-graph", action="store_true", help="launch every step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--quick", action="store_true", help="skip the RMSD / configs[4] / dense / configs[1] side measurements")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "rmsd":
        run_rmsd_only(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
