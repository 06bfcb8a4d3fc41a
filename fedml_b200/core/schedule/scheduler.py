"""Workload-aware assignment of simulated clients to GPUs.

Semantics follow the reference's unreferenced ``core/schedule/scheduler.py:4-183`` (the
"Resource Scheduling" its NCCL-simulator README alludes to): workloads ``x`` (e.g. sample
counts), per-resource cost factors ``y`` (seconds per unit of work; lower = faster GPU) and
per-resource memory caps ``m``.  The objective is to minimise the maximum per-resource cost.
``mode 0``: jobs on a resource run serially (cost adds up).  ``mode 1``: jobs may also be
packed to run in parallel "bunches" as long as the bunch fits in memory.

The reference explores the full assignment tree recursively, which is exponential and
overflows the Python stack at ~100 clients.  Here the exact search is an explicit best-first
branch-and-bound with an expansion budget; beyond it (or for > ``exact_limit`` jobs) we fall
back to LPT (longest processing time first) list scheduling, which is within 4/3 of optimal
for identical machines and is what the NCCL simulator uses every round.
"""
from __future__ import annotations

import heapq
import itertools
from typing import Dict, List, Sequence

import numpy as np


def lpt_assign(workloads: Sequence[float], speeds: Sequence[float], memory: Sequence[float] = None,
               max_count: int = 0) -> List[List[int]]:
    """Greedy LPT: biggest job first onto the resource whose finishing cost stays smallest.
    ``max_count`` > 0 caps the number of jobs per resource (the client-batched executor sizes its stacks by it)."""
    x = np.asarray(workloads, dtype=np.float64)
    y = np.asarray(speeds, dtype=np.float64)
    order = np.argsort(-x, kind="stable")
    load = np.zeros(len(y))
    count = np.zeros(len(y), dtype=np.int64)
    out: List[List[int]] = [[] for _ in y]
    for j in order:
        cost = load + y * x[j]
        if max_count > 0:
            cost = np.where(count >= max_count, np.inf, cost)
        # tie-break on job count so equal workloads spread evenly (keeps lockstep batches square)
        r = int(np.lexsort((count, cost))[0])
        out[r].append(int(j))
        load[r] = cost[r]
        count[r] += 1
    return out


class scheduler:  # noqa: N801 - reference class name
    def __init__(self, workloads, constraints, memory, exact_limit: int = 12, max_expansions: int = 200_000):
        self.workloads = np.asarray(workloads)
        self.x_sorted_index = np.argsort(self.workloads, kind="stable")[::-1]
        self.x = self.workloads[self.x_sorted_index]
        self.y = np.asarray(constraints, dtype=np.float64)
        self.m = np.asarray(memory, dtype=np.float64)
        self.len_x, self.len_y = len(self.x), len(self.y)
        self.exact_limit, self.max_expansions = exact_limit, max_expansions
        self.last_cost = None

    # -- exact search ------------------------------------------------------
    def _search(self, mode: int):
        tie = itertools.count()
        start = (0.0, next(tie), 0, tuple([-1] * self.len_x), tuple([0.0] * self.len_y), tuple([0.0] * self.len_y))
        heap = [start]
        expansions = 0
        while heap:
            cost, _, depth, amap, costs, mem = heapq.heappop(heap)
            if depth == self.len_x:
                return amap, cost
            expansions += 1
            if expansions > self.max_expansions:
                return None, None
            w = float(self.x[depth])
            for r in range(self.len_y):
                variants = []
                if mode == 1:
                    # parallel with what is already resident on r
                    c_par = list(costs); c_par[r] = max(self.y[r] * w, c_par[r])
                    m_par = list(mem); m_par[r] += w
                    variants.append((c_par, m_par))
                    c_ser = list(costs); c_ser[r] += self.y[r] * w
                    m_ser = list(mem); m_ser[r] = w
                    variants.append((c_ser, m_ser))
                else:
                    c_ser = list(costs); c_ser[r] += self.y[r] * w
                    variants.append((c_ser, list(mem)))
                for c_new, m_new in variants:
                    if mode == 1:
                        if m_new[r] > self.m[r]:
                            continue
                    elif max(c_new) > self.m[int(np.argmax(c_new))]:
                        continue  # reference quirk: serial mode compares cost to the memory cap
                    nm = list(amap); nm[depth] = r
                    heapq.heappush(heap, (max(c_new), next(tie), depth + 1, tuple(nm), tuple(c_new), tuple(m_new)))
        return None, None

    def DP_schedule(self, mode: int = 0) -> List[Dict[int, List[int]]]:  # noqa: N802
        amap = None
        if self.len_x <= self.exact_limit:
            amap, cost = self._search(mode)
            self.last_cost = cost
        if amap is None:
            groups = lpt_assign(self.x, self.y, self.m)
            amap = [-1] * self.len_x
            for r, js in enumerate(groups):
                for j in js:
                    amap[j] = r
            loads = [sum(self.y[r] * self.x[j] for j in js) for r, js in enumerate(groups)]
            self.last_cost = max(loads) if loads else 0.0
        schedules = [[int(self.x_sorted_index[j]) for j in range(self.len_x) if amap[j] == r]
                     for r in range(self.len_y)]
        out = []
        for r, seq in enumerate(schedules):
            plan: Dict[int, List[int]] = {}
            if mode == 1:
                jobs, foot = [], 0.0
                for job in seq:
                    w = float(self.workloads[job])
                    if jobs and foot + w > self.m[r]:
                        plan[len(plan)] = jobs
                        jobs, foot = [], 0.0
                    jobs.append(job); foot += w
                if jobs:
                    plan[len(plan)] = jobs
            else:
                for job in seq:
                    plan[len(plan)] = [job]
            out.append(plan)
        return out


def schedule_clients(sample_counts: Sequence[int], n_gpus: int, speeds: Sequence[float] = None) -> List[List[int]]:
    """Round-path entry point: positions (0..K-1) of the sampled clients per GPU rank."""
    speeds = np.ones(n_gpus) if speeds is None else np.asarray(speeds, dtype=np.float64)
    cap = (len(sample_counts) + n_gpus - 1) // n_gpus
    return lpt_assign(sample_counts, speeds, max_count=cap)


# ``fedml.core.schedule.scheduler`` is both this module and (re-exported by the package) the class of the same name, as in the
# reference.  Importing the sub-module rebinds the package attribute from the class to the module, so the module is made
# callable: ``fedml.core.schedule.scheduler(...)`` builds a scheduler whichever of the two the name currently points at.
import sys as _sys
import types as _types


class _CallableModule(_types.ModuleType):
    def __call__(self, *args, **kwargs):
        return scheduler(*args, **kwargs)


_sys.modules[__name__].__class__ = _CallableModule
