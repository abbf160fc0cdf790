"""The handful of DEAP 1.2 pieces the GaudiMM GA loop uses (gaudi/algorithms.py:54-202,
gaudi/cli/gaudi_run.py:69-134, gaudi/genes/torsion.py:173-195), restated from DEAP's
published algorithms because DEAP is not installable here.  All of it is host-side GA
bookkeeping ("next" row 1); none of it is on the CUDA path.

Attribution: ``cx_simulated_binary_bounded``, ``mut_polynomial_bounded``, ``cx_two_point``,
``var_or``, ``sort_nondominated``, ``assign_crowding_dist``, ``sel_nsga2``, ``HallOfFame`` and
``ParetoFront`` follow the corresponding functions of DEAP (Distributed Evolutionary Algorithms in
Python, https://github.com/DEAP/deap, (c) the DEAP authors, GNU LGPL v3), operation for operation,
so that a GA run varies and selects alleles exactly as the reference's DEAP calls do.
"""
from __future__ import annotations

import random
from copy import deepcopy
from operator import attrgetter


class Fitness(object):
    """``deap.base.Fitness`` as subclassed at gaudi/base.py:371-382: ``weights`` per instance,
    ``values`` are the raw scores, ``wvalues = values * weights`` drive comparisons."""

    def __init__(self, weights):
        self.weights = tuple(weights)
        self.wvalues = ()

    @property
    def values(self):
        return tuple(w / wt for w, wt in zip(self.wvalues, self.weights))

    @values.setter
    def values(self, values):
        self.wvalues = tuple(v * w for v, w in zip(values, self.weights))

    @values.deleter
    def values(self):
        self.wvalues = ()

    @property
    def valid(self):
        return len(self.wvalues) != 0

    def dominates(self, other):
        not_equal = False
        for a, b in zip(self.wvalues, other.wvalues):
            if a > b:
                not_equal = True
            elif a < b:
                return False
        return not_equal

    def __le__(self, other):
        return self.wvalues <= other.wvalues

    def __lt__(self, other):
        return self.wvalues < other.wvalues

    def __ge__(self, other):
        return not self.__lt__(other)

    def __gt__(self, other):
        return not self.__le__(other)

    def __eq__(self, other):
        return self.wvalues == other.wvalues

    def __ne__(self, other):
        return not self.__eq__(other)

    def __hash__(self):
        return hash(self.wvalues)

    def __deepcopy__(self, memo):
        new = self.__class__(self.weights)
        new.wvalues = self.wvalues + ()
        return new


class Toolbox(object):
    """``deap.base.Toolbox``: named aliases with frozen arguments; default ``map`` is the
    builtin and ``clone`` is ``deepcopy``."""

    def __init__(self):
        self.register("clone", deepcopy)
        self.register("map", lambda fn, *its: list(map(fn, *its)))

    def register(self, alias, function, *args, **kargs):
        if args or kargs:
            def pfunc(*a, **k):
                kw = dict(kargs)
                kw.update(k)
                return function(*(args + a), **kw)
            pfunc.__name__ = alias
        else:
            pfunc = function
        setattr(self, alias, pfunc)

    def unregister(self, alias):
        delattr(self, alias)


def var_or(population, toolbox, lambda_, cxpb, mutpb):
    """``deap.algorithms.varOr``: each child comes from crossover (first child kept), mutation
    or plain reproduction."""
    assert (cxpb + mutpb) <= 1.0
    offspring = []
    for _ in range(lambda_):
        op = random.random()
        if op < cxpb:
            ind1, ind2 = map(toolbox.clone, random.sample(population, 2))
            ind1, ind2 = toolbox.mate(ind1, ind2)
            del ind1.fitness.values
            offspring.append(ind1)
        elif op < cxpb + mutpb:
            ind = toolbox.clone(random.choice(population))
            ind, = toolbox.mutate(ind)
            del ind.fitness.values
            offspring.append(ind)
        else:
            offspring.append(random.choice(population))
    return offspring


# ---- NSGA-II --------------------------------------------------------------------------
def sort_nondominated(individuals, k, first_front_only=False):
    if k == 0:
        return []
    map_fit = {}
    for ind in individuals:
        map_fit.setdefault(ind.fitness, []).append(ind)
    fits = list(map_fit.keys())
    current, nxt = [], []
    dominating = {f: 0 for f in fits}
    dominated = {f: [] for f in fits}
    for i, fi in enumerate(fits):
        for fj in fits[i + 1:]:
            if fi.dominates(fj):
                dominating[fj] += 1
                dominated[fi].append(fj)
            elif fj.dominates(fi):
                dominating[fi] += 1
                dominated[fj].append(fi)
        if dominating[fi] == 0:
            current.append(fi)
    fronts = [[]]
    for f in current:
        fronts[-1].extend(map_fit[f])
    pareto_sorted = len(fronts[-1])
    if not first_front_only:
        n = min(len(individuals), k)
        while pareto_sorted < n:
            fronts.append([])
            for fp in current:
                for fd in dominated[fp]:
                    dominating[fd] -= 1
                    if dominating[fd] == 0:
                        nxt.append(fd)
                        pareto_sorted += len(map_fit[fd])
                        fronts[-1].extend(map_fit[fd])
            current, nxt = nxt, []
    return fronts


def assign_crowding_dist(individuals):
    if not individuals:
        return
    distances = [0.0] * len(individuals)
    crowd = [(ind.fitness.values, i) for i, ind in enumerate(individuals)]
    nobj = len(individuals[0].fitness.values)
    for i in range(nobj):
        crowd.sort(key=lambda e: e[0][i])
        distances[crowd[0][1]] = float("inf")
        distances[crowd[-1][1]] = float("inf")
        if crowd[-1][0][i] == crowd[0][0][i]:
            continue
        norm = nobj * float(crowd[-1][0][i] - crowd[0][0][i])
        for prev, cur, nxt in zip(crowd[:-2], crowd[1:-1], crowd[2:]):
            distances[cur[1]] += (nxt[0][i] - prev[0][i]) / norm
    for i, d in enumerate(distances):
        individuals[i].fitness.crowding_dist = d


def sel_nsga2(individuals, k):
    """``deap.tools.selNSGA2`` (registered as ``select`` at gaudi/cli/gaudi_run.py:95)."""
    fronts = sort_nondominated(individuals, k)
    for front in fronts:
        assign_crowding_dist(front)
    chosen = [ind for front in fronts[:-1] for ind in front]
    k = k - len(chosen)
    if k > 0:
        last = sorted(fronts[-1], key=attrgetter("fitness.crowding_dist"), reverse=True)
        chosen.extend(last[:k])
    return chosen


# ---- halls of fame ----------------------------------------------------------------------
class HallOfFame(object):
    def __init__(self, maxsize, similar=None):
        self.maxsize = maxsize
        self.items = []
        self.similar = similar or (lambda a, b: a is b)

    def update(self, population):
        for ind in population:
            if len(self.items) < self.maxsize or ind.fitness > self.items[-1].fitness:
                if not any(self.similar(ind, h) for h in self.items):
                    if len(self.items) >= self.maxsize:
                        self.items.pop()
                    self.items.append(deepcopy(ind))
                    self.items.sort(key=attrgetter("fitness"), reverse=True)

    def clear(self):
        del self.items[:]

    def __len__(self):
        return len(self.items)

    def __getitem__(self, i):
        return self.items[i]

    def __iter__(self):
        return iter(self.items)


class ParetoFront(HallOfFame):
    """``deap.tools.ParetoFront`` with a similarity test (gaudi/cli/gaudi_run.py:108-109)."""

    def __init__(self, similar=None):
        HallOfFame.__init__(self, None, similar)

    def update(self, population):
        for ind in population:
            is_dominated, has_twin, to_remove = False, False, []
            for i, hofer in enumerate(self.items):
                if hofer.fitness.dominates(ind.fitness):
                    is_dominated = True
                    break
                elif ind.fitness.dominates(hofer.fitness):
                    to_remove.append(i)
                elif ind.fitness == hofer.fitness and self.similar(ind, hofer):
                    has_twin = True
                    break
            for i in reversed(to_remove):
                del self.items[i]
            if not is_dominated and not has_twin:
                # DEAP's HallOfFame.insert keeps the items sorted best first; the reference writes its
                # results in that order, so result numbering follows it
                self.items.append(deepcopy(ind))
                self.items.sort(key=attrgetter("fitness"), reverse=True)


# ---- variation operators ----------------------------------------------------------------
def cx_simulated_binary_bounded(ind1, ind2, eta, low, up):
    """``deap.tools.cxSimulatedBinaryBounded`` (torsion.py:174-180)."""
    size = min(len(ind1), len(ind2))
    for i in range(size):
        if random.random() <= 0.5:
            if abs(ind1[i] - ind2[i]) > 1e-14:
                x1, x2 = min(ind1[i], ind2[i]), max(ind1[i], ind2[i])
                rand = random.random()
                beta = 1.0 + (2.0 * (x1 - low) / (x2 - x1))
                alpha = 2.0 - beta ** -(eta + 1)
                if rand <= 1.0 / alpha:
                    beta_q = (rand * alpha) ** (1.0 / (eta + 1))
                else:
                    beta_q = (1.0 / (2.0 - rand * alpha)) ** (1.0 / (eta + 1))
                c1 = 0.5 * (x1 + x2 - beta_q * (x2 - x1))
                beta = 1.0 + (2.0 * (up - x2) / (x2 - x1))
                alpha = 2.0 - beta ** -(eta + 1)
                if rand <= 1.0 / alpha:
                    beta_q = (rand * alpha) ** (1.0 / (eta + 1))
                else:
                    beta_q = (1.0 / (2.0 - rand * alpha)) ** (1.0 / (eta + 1))
                c2 = 0.5 * (x1 + x2 + beta_q * (x2 - x1))
                c1, c2 = min(max(c1, low), up), min(max(c2, low), up)
                if random.random() <= 0.5:
                    ind1[i], ind2[i] = c2, c1
                else:
                    ind1[i], ind2[i] = c1, c2
    return ind1, ind2


def mut_polynomial_bounded(individual, eta, low, up, indpb):
    """``deap.tools.mutPolynomialBounded`` (torsion.py:186-193)."""
    for i in range(len(individual)):
        if random.random() <= indpb:
            x = individual[i]
            delta_1 = (x - low) / (up - low)
            delta_2 = (up - x) / (up - low)
            rand = random.random()
            mut_pow = 1.0 / (eta + 1.0)
            if rand < 0.5:
                xy = 1.0 - delta_1
                val = 2.0 * rand + (1.0 - 2.0 * rand) * xy ** (eta + 1)
                delta_q = val ** mut_pow - 1.0
            else:
                xy = 1.0 - delta_2
                val = 2.0 * (1.0 - rand) + 2.0 * (rand - 0.5) * xy ** (eta + 1)
                delta_q = 1.0 - val ** mut_pow
            x = x + delta_q * (up - low)
            individual[i] = min(max(x, low), up)
    return individual,


def cx_two_point(ind1, ind2):
    """``deap.tools.cxTwoPoint`` (rotamers.py:160)."""
    size = min(len(ind1), len(ind2))
    if size < 2:
        return ind1, ind2
    p1 = random.randint(1, size)
    p2 = random.randint(1, size - 1)
    if p2 >= p1:
        p2 += 1
    else:
        p1, p2 = p2, p1
    ind1[p1:p2], ind2[p1:p2] = ind2[p1:p2], ind1[p1:p2]
    return ind1, ind2
