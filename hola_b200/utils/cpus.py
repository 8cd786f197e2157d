"""CPU partitioning for the load test.

SURVEY.md 6.3: the load generator and the server share one machine, so they are pinned to
disjoint core sets and the split is recorded with the result; otherwise the number measures the
scheduler's mood, not the service.
"""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple


def available_cpus() -> List[int]:
    return sorted(os.sched_getaffinity(0))


def _siblings(cpu: int) -> List[int]:
    path = "/sys/devices/system/cpu/cpu%d/topology/thread_siblings_list" % cpu
    try:
        text = open(path).read().strip()
    except OSError:
        return [cpu]
    out: List[int] = []
    for part in text.split(","):
        if "-" in part:
            a, b = part.split("-")
            out.extend(range(int(a), int(b) + 1))
        elif part:
            out.append(int(part))
    return out or [cpu]


def split_cpus(cpus: Sequence[int], server_share: float = 0.5, max_total: int = 0) -> Tuple[List[int], List[int]]:
    """Splits ``cpus`` into (server, load generator) sets.

    Hyper-thread siblings are kept on the same side where the topology is visible, so the two
    programs do not fight over one physical core. ``max_total`` (if > 0) caps how many CPUs are
    used overall - beyond a few dozen cores 64 connections cannot keep more threads busy anyway.
    """
    cpus = list(cpus)
    if len(cpus) < 2:
        return cpus, cpus
    # Group by physical core.
    seen, cores = set(), []
    allowed = set(cpus)
    for c in cpus:
        if c in seen:
            continue
        group = [s for s in _siblings(c) if s in allowed] or [c]
        for s in group:
            seen.add(s)
        cores.append(group)
    if max_total > 0:
        kept, total = [], 0
        for g in cores:
            if total + len(g) > max_total and kept:
                break
            kept.append(g)
            total += len(g)
        cores = kept
    n_server = max(1, min(len(cores) - 1, round(len(cores) * server_share))) if len(cores) > 1 else 1
    server = [c for g in cores[:n_server] for c in g]
    load = [c for g in cores[n_server:] for c in g] or server
    return server, load
