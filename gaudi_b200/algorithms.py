"""The (mu + lambda) loop of gaudi/algorithms.py:54-202 with the two seams of the B200 path
filled in: ``toolbox.map`` is a batched evaluate (one GPU call per generation instead of a
serial Python ``map``), and the O((mu+lambda)^2) ``toolbox.similarity`` calls of the niching
pass (algorithms.py:139-155) read a bitmask computed by one ``gmm_rmsd_matrix`` call.  The
greedy, order-dependent niching pass itself, hall of fame and NSGA-II selection run on the host
exactly as in the reference."""
from __future__ import annotations

import logging
import random
from time import time

from . import deaplite

logger = logging.getLogger(__name__)


class BatchedMap(object):
    """Drop-in for ``toolbox.map``: ``map(toolbox.evaluate, individuals)`` becomes one batched
    GPU evaluation.  Any other mapped function falls back to the builtin."""

    def __init__(self, environment, evaluate_alias=None):
        self.environment = environment
        self.evaluate_alias = evaluate_alias

    def __call__(self, fn, individuals):
        individuals = list(individuals)
        if self.evaluate_alias is None or fn is self.evaluate_alias:
            return self.environment.evaluate_population(individuals)
        return list(map(fn, individuals))


def niching(candidates, similar):
    """gaudi/algorithms.py:139-155 verbatim in behaviour: visit in order; drop a candidate at the
    first similar kept individual that is at least as fit, otherwise evict the similar, worse ones."""
    new_population = []
    for i in candidates:
        to_delete = []
        for n_i in new_population:
            if similar(i, n_i):
                if i.fitness <= n_i.fitness:
                    break
                else:
                    to_delete.append(new_population.index(n_i))
        else:
            if to_delete:
                new_population = [k for j, k in enumerate(new_population) if j not in to_delete]
            new_population.append(i)
    return new_population


def ea_mu_plus_lambda(population, toolbox, mu, lambda_, cxpb, mutpb, ngen, cfg=None, stats=None,
                      halloffame=None, verbose=True, environment=None):
    """Returns ``(population, logbook)``; ``logbook`` is a list of dicts with the reference's
    columns ``gen, nevals, speed`` (ev/s defined as in algorithms.py:168-171)."""
    logbook = []
    t0 = time()
    invalid = [ind for ind in population if not ind.fitness.valid]
    for ind, fit in zip(invalid, toolbox.map(toolbox.evaluate, invalid)):
        ind.fitness.values = fit
    population = toolbox.select(population, len(population))
    if halloffame is not None:
        halloffame.update(population)
    nevals = len(invalid)
    logbook.append({"gen": 0, "nevals": nevals, "speed": nevals / max(time() - t0, 1e-9)})
    batch = environment.batch if environment is not None else None
    sim_args = cfg.similarity.args if cfg is not None else None
    for gen in range(1, ngen + 1):
        offspring = deaplite.var_or(population, toolbox, lambda_, cxpb, mutpb)
        t1 = time()
        invalid = [ind for ind in offspring if not ind.fitness.valid]
        for ind, fit in zip(invalid, toolbox.map(toolbox.evaluate, invalid)):
            ind.fitness.values = fit
        candidates = population + offspring
        if batch is not None and sim_args is not None and getattr(toolbox, "array_niching", False):
            # one GPU call per generation; the greedy pass runs over the partner lists it leaves
            # (gmm_niching_host), same result as the pairwise loop below
            import numpy as np
            from . import tensor_ga
            lists = batch.similarity_lists(candidates, sim_args[0], sim_args[1])
            w = np.array([ind.fitness.wvalues for ind in candidates], dtype=np.float64)
            new_population = [candidates[i] for i in tensor_ga.niching(lists, w)]
        else:
            if batch is not None and sim_args is not None:
                batch.similarity_matrix(candidates, sim_args[0], sim_args[1])
            new_population = niching(candidates, toolbox.similarity)
        while len(new_population) < mu:
            new_population.append(random.choice(candidates))
        if halloffame is not None:
            halloffame.clear()
            halloffame.update(new_population)
        population[:] = toolbox.select(new_population, mu)
        nevals = len(invalid)
        t2 = time()
        logbook.append({"gen": gen, "nevals": nevals, "speed": nevals / max(t2 - t1, 1e-9)})
        if verbose:
            logger.log(100, "gen %d  nevals %d  %.2f ev/s", gen, nevals, logbook[-1]["speed"])
        # gaudi/algorithms.py:197-201: a copy of the elite every check_every generations, in case
        # the run is stopped later
        if halloffame and cfg is not None and cfg.output.check_every and not gen % cfg.output.check_every:
            try:
                dump_population(halloffame, cfg, subdir="check{}".format(gen))
            except Exception:
                logger.warning("Could not write checkpoint for gen #%s", gen)
    return population, logbook


def dump_population(population, cfg, subdir=None):
    """gaudi/algorithms.py:205-223: ``Individual.write`` for every member (``<name>_<i>.zip``) plus
    the ``<name>.gaudi-output`` yaml index of file name -> scores.  Returns the index path."""
    import os
    from time import strftime
    import yaml
    results = {"GAUDI.objectives": ["{} ({})".format(obj.name, obj.module) for obj in cfg.objectives],
               "GAUDI.results": {}}
    basepath = cfg.output.path
    if subdir is not None:
        basepath = os.path.join(cfg.output.path, subdir)
        os.makedirs(basepath, exist_ok=True)
    for i, ind in enumerate(population):
        filename = ind.write(i, path=basepath)
        results["GAUDI.results"][os.path.basename(filename)] = [float(v) for v in ind.fitness.values]
    gaudi_output = os.path.join(basepath, cfg.output.name + ".gaudi-output")
    with open(gaudi_output, "w+") as out:
        out.write("# Generated by gaudi_b200 (GaudiMM hot path on B200) on {}\n\n".format(strftime("%Y-%m-%d %H:%M:%S")))
        out.write(yaml.safe_dump(results, default_flow_style=False))
    return gaudi_output


def launch(cfg, device=0, seed=None, group=None):
    """gaudi/cli/gaudi_run.py:69-134 for the B200 path: build population and environment,
    register the DEAP-style aliases plus the batched ``map``, run the loop.

    Several GPUs: start one process per GPU (``torchrun``), initialise ``torch.distributed`` and call
    ``launch`` in every process with its own ``device``.  Rank 0 runs the GA and returns
    ``(population, logbook, elite)``; the other ranks evaluate their share of every
    ``toolbox.map(toolbox.evaluate, ...)`` call and return ``(None, None, None)`` when rank 0 is done.
    Results are identical to a single-GPU run with the same seed."""
    from . import base, parallel, plugin
    if seed is not None:
        random.seed(seed)
    plugin.import_plugins(*cfg.genes)
    plugin.import_plugins(*cfg.objectives)
    toolbox = deaplite.Toolbox()
    world, rank = parallel._world(group) if group is not False else (1, 0)
    environment = base.Environment(cfg, device=device, group=group)
    cfg._environment = environment
    if rank != 0:
        environment.attach(base.MolecularIndividual(cfg))
        environment.batch.enable_sharding(cfg.ga.population + int(cfg.ga.lambda_ * cfg.ga.population))
        environment.batch.serve()
        environment.batch.close()
        return None, None, None
    population = [base.MolecularIndividual(cfg) for _ in range(cfg.ga.population)]
    environment.attach(population[0])
    environment.batch.enable_sharding(cfg.ga.population + int(cfg.ga.lambda_ * cfg.ga.population))
    toolbox.register("evaluate", environment.evaluate)
    toolbox.register("mate", lambda a, b: a.mate(b))
    toolbox.register("mutate", lambda ind, indpb: ind.mutate(indpb), indpb=cfg.ga.mut_indpb)
    toolbox.register("similarity", lambda a, b: a.similar(b))
    if cfg.similarity.module == "gaudi_b200.similarity.rmsd":
        # the stock similarity: niching and selection work on arrays (partner lists from the
        # device, fronts by gmm_fronts2_host / gmm_nondominated_fronts) instead of O(N^2)
        # Python-level calls; any other similarity function keeps the reference's pairwise loop
        toolbox.array_niching = True
        from . import tensor_ga
        import numpy as np
        tensor_ga.set_fronts_backend(environment.batch.engines[0][0].nondominated_fronts)

        def select(individuals, k):
            values = np.array([ind.fitness.values for ind in individuals], dtype=np.float64)
            return [individuals[i] for i in tensor_ga.sel_nsga2(values, np.asarray(cfg.weights, float), k)]
        toolbox.register("select", select)
    else:
        toolbox.register("select", deaplite.sel_nsga2)
    toolbox.register("map", BatchedMap(environment, toolbox.evaluate))
    if cfg.output.pareto:
        elite = deaplite.ParetoFront(toolbox.similarity)
    else:
        elite = deaplite.HallOfFame(int(cfg.ga.population * cfg.ga.mu), similar=toolbox.similarity)
    mu = int(cfg.ga.mu * cfg.ga.population)
    lambda_ = int(cfg.ga.lambda_ * cfg.ga.population)
    population, log = ea_mu_plus_lambda(population, toolbox, mu=mu, lambda_=lambda_, cxpb=cfg.ga.cx_pb,
                                        mutpb=cfg.ga.mut_pb, ngen=cfg.ga.generations, cfg=cfg,
                                        halloffame=elite, verbose=cfg.output.verbose, environment=environment)
    environment.batch.shutdown()
    return population, log, elite


def run(cfg, device=0, seed=None, group=None):
    """What ``gaudi run`` does after parsing its input (gaudi/cli/gaudi_run.py:137-178): ``launch``,
    then write the elite with ``dump_population``.  Returns ``(population, logbook, elite, index
    path)``; ranks other than 0 of a multi-GPU job return ``(None, None, None, None)``."""
    import os
    population, log, elite = launch(cfg, device=device, seed=seed, group=group)
    if population is None:
        return None, None, None, None
    os.makedirs(cfg.output.path, exist_ok=True)
    index = dump_population(elite, cfg)
    return population, log, elite, index
