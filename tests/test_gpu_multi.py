"""Several ranks (one process per rank, ``torch.distributed``) against one process, on the GPU.

With two or more GPUs every rank takes its own; on a one-GPU box both ranks share cuda:0 — the
peer mappings (CUDA IPC), the epoch flags and the shared host table are the same code either
way, only the wire under the peer stores differs.  The process group is gloo (rendezvous and
handle exchange only), so that two ranks may share a GPU."""
import os
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _init(rank, world, port):
    import sys
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    return torch, dist, dev


def _exchange_worker(rank, world, port, out):
    torch, dist, dev = _init(rank, world, port)
    from gaudi_b200 import model as M, parallel, synth
    from gaudi_b200.engine import Engine
    params = M.ContactParams()
    lig = synth.make_ligand(40, 10, seed=1)
    rec = synth.make_receptor(3000, seed=2, params=params)
    eng = Engine(lig, rec, params, device=dev)
    pop = 5001                                                      # ragged over the ranks
    results = {}
    # "nccl" = the collective variant (all_gather_into_tensor; under this test's gloo group it is staged
    # through the host), kept as the cross-check of the peer-store exchange
    sh = parallel.ShardedEvaluator(eng, pop, exchange="nccl")
    tors, xf = synth.make_alleles(pop, 10, seed=2)
    s, c = sh.evaluate(torch.as_tensor(sh.shard(tors)).to(eng.torch_device), torch.as_tensor(sh.shard(xf)).to(eng.torch_device))
    want_s, want_c, _ = eng.evaluate(tors, xf)
    results["collective"] = bool(torch.equal(s, want_s) and torch.equal(c, want_c))
    sh.close()
    for kind in ("peer",):
        sh = parallel.ShardedEvaluator(eng, pop, exchange=kind)
        for step, seed in enumerate((3, 4, 5)):                     # three epochs: both buffers, then reuse
            tors, xf = synth.make_alleles(pop, 10, seed=seed)
            t = torch.as_tensor(sh.shard(tors)).to(eng.torch_device)
            x = torch.as_tensor(sh.shard(xf)).to(eng.torch_device)
            s, c = sh.evaluate(t, x)
            torch.cuda.synchronize()
            want_s, want_c, _ = eng.evaluate(tors, xf)
            results[(kind, step)] = bool(torch.equal(s, want_s) and torch.equal(c, want_c))
        epoch, timed_out, s_last, _ = sh.exchange.state()
        results[(kind, "state")] = (epoch == 3 and not timed_out and bool(torch.equal(s_last[:pop], want_s)))
        # the same step replayed from a CUDA graph: no collective, no host work inside
        stream = torch.cuda.Stream(device=eng.torch_device)
        stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(stream):
            sh.evaluate(t, x)
        torch.cuda.current_stream().wait_stream(stream)
        torch.cuda.synchronize()
        dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            sh.evaluate(t, x)
        for _ in range(3):
            g.replay()
        epoch, timed_out, s_last, c_last = sh.exchange.state()
        # 3 eager steps + 1 warm-up on the side stream + 3 replays (the capture itself runs nothing)
        results[(kind, "graph")] = (epoch == 7 and not timed_out and bool(torch.equal(s_last[:pop], want_s))
                                    and bool(torch.equal(c_last[:pop], want_c)))
        dist.barrier()
        sh.close()
    # row-block sharded similarity: only the partner lists travel
    conf = synth.make_conformers(1500, 60, seed=4).astype(np.float32)
    b = parallel.shard_bounds(1500, world)
    mine = torch.as_tensor(conf[b[rank]:b[rank + 1]]).to(eng.torch_device)
    ptr, idx = parallel.sharded_similar_lists(eng, mine, [b[r + 1] - b[r] for r in range(world)], threshold=2.5)
    _, bits = eng.rmsd_matrix(conf, threshold=2.5, values=False)
    ptr1, idx1 = eng.similar_lists(bits, 1500)
    results["lists"] = bool(np.array_equal(ptr.cpu().numpy(), ptr1) and np.array_equal(idx.cpu().numpy(), idx1)
                            and len(idx1) > 0)
    # the same through the preallocated generational path (one host synchronisation per step), with a
    # capacity too small for the first step so that the retry is exercised too
    sim = parallel.ShardedSimilarity(eng, [b[r + 1] - b[r] for r in range(world)], 60, threshold=2.5, cap=16)
    for _ in range(2):
        ptr2, idx2 = sim.step(mine)
        results["lists_prealloc"] = bool(np.array_equal(ptr2, ptr1) and np.array_equal(idx2, idx1)) and results.get("lists_prealloc", True)
    # the lists on rank 0 alone (where the niching pass runs): the other ranks skip the copy of the entries
    ptr3, idx3 = sim.step(mine, root_only=True)
    if rank == 0:
        results["lists_root_only"] = bool(np.array_equal(ptr3, ptr1) and np.array_equal(idx3, idx1))
    else:
        assert ptr3 is None and idx3 is None
    dist.barrier()
    # host-resident population: shared page-locked table, each rank uploads and downloads its rows
    hs = parallel.HostShardedEvaluator(eng, 6000, want_xyz=True)
    if rank == 0:
        ok = True
        for seed, n in ((7, 6000), (8, 1234), (9, 1)):
            tors, xf = synth.make_alleles(n, 10, seed=seed)
            hs.tors[:n] = tors; hs.xforms[:n] = xf
            hs.scores[:] = np.nan
            hs.evaluate(n)
            want_s, want_c, want_x = eng.evaluate(tors, xf, want_xyz=True)
            ok = ok and np.array_equal(hs.scores[:n], want_s.cpu().numpy()) \
                and np.array_equal(hs.counts[:n], want_c.cpu().numpy()) \
                and np.array_equal(hs.xyz[:n], want_x.cpu().numpy())
        results["host"] = bool(ok)
        hs.shutdown()
    else:
        hs.serve()
    hs.close()
    dist.barrier()
    if rank == 0:
        with open(out, "w") as fh:
            fh.write(repr(sorted((str(k), v) for k, v in results.items())))
    eng.close()
    dist.destroy_process_group()


def _spawn(fn, world, tmp_path, name):
    import torch
    import torch.multiprocessing as mp
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    out = str(tmp_path / name)
    port = 33500 + random.randint(0, 2000)
    mp.spawn(fn, args=(world, port, out), nprocs=world, join=True)
    return out


def test_fitness_exchange_and_host_table_two_ranks(built_lib, tmp_path):
    """gmm_exchange_* (peer stores + epoch flags) gathers exactly the rows a single evaluate
    produces, eagerly and replayed from a CUDA graph; the shared host table does the same for
    host-resident populations."""
    out = _spawn(_exchange_worker, 2, tmp_path, "ex.txt")
    res = dict(eval(open(out).read()))
    assert res and all(res.values()), res


def _launch_worker(rank, world, port, out):
    torch, dist, dev = _init(rank, world, port)
    import test_gpu_plugins as TP
    from gaudi_b200 import algorithms, tensor_ga
    cfg = TP.job(60, 4)
    pop, log, elite = algorithms.launch(cfg, device=dev, seed=11)
    dist.barrier()
    cfg2 = TP.job(300, 3)
    tpop, tlog, batch = tensor_ga.launch(cfg2, device=dev, seed=12)
    dist.barrier()
    if rank == 0:
        fit = np.array([ind.fitness.values for ind in pop])
        alle = np.array([ind.genes["Torsion"].allele for ind in pop])
        np.savez(out, fit=fit, alle=alle, tfit=tpop["fitness"], ttors=tpop["torsions"])
    else:
        assert pop is None and tpop is None
    dist.destroy_process_group()


def test_two_rank_launch_equals_single_process_launch(built_lib, tmp_path):
    """``algorithms.launch`` / ``tensor_ga.launch`` with the population's evaluation sharded over
    two ranks evolve exactly the population a single process evolves from the same seed."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import test_gpu_plugins as TP
    from gaudi_b200 import algorithms, tensor_ga
    out = _spawn(_launch_worker, 2, tmp_path, "launch.npz")
    got = np.load(out)
    pop, _, _ = algorithms.launch(TP.job(60, 4), seed=11, group=False)
    fit = np.array([ind.fitness.values for ind in pop])
    alle = np.array([ind.genes["Torsion"].allele for ind in pop])
    assert np.array_equal(got["fit"], fit) and np.array_equal(got["alle"], alle)
    tpop, _, _ = tensor_ga.launch(TP.job(300, 3), seed=12, group=False)
    assert np.array_equal(got["tfit"], tpop["fitness"]) and np.array_equal(got["ttors"], tpop["torsions"])
