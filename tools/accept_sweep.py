#!/usr/bin/env python
"""Sweep of connection-placement modes x worker counts (VERDICT r1 "weak" item 1).

For every (workers, mode) pair: start `hola` pinned to its CPUs, drive 64 keep-alive connections
with `hola-load` pinned to the other CPUs, and record requests/sec, latency and how evenly the
workers shared the work (max/min requests per worker, and the per-worker connection counts).

    python tools/accept_sweep.py [--workers 4,8,16] [--seconds 8] [--repeat 3] [--out FILE]

The reference gets an even spread from Undertow's acceptor (/root/reference/pom.xml:122-129).
"""
from __future__ import annotations

import argparse
import json
import statistics
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

from hola_b200 import build, loadgen  # noqa: E402
from hola_b200.server import HolaServer  # noqa: E402
from hola_b200.utils import available_cpus  # noqa: E402


def one_run(workers: int, mode: str, seconds: int, server_cpus, load_cpus, connections: int, path: str, extra):
    args = ["--cpus", ",".join(map(str, server_cpus)), "--accept-mode", mode] + list(extra)
    with HolaServer(args=args, env={"HOSTNAME": None, "ZIPKIN_SERVER_URL": None}, threads=workers) as srv:
        before = srv.stats()
        rep = loadgen.run(srv.base_url + path, connections=connections, duration=seconds, warmup=1,
                          threads=min(len(load_cpus), connections), cpus=load_cpus)
        after = srv.stats()
    req = [a["requests"] - b["requests"] for a, b in zip(after["workers"], before["workers"])]
    cpu = [a["cpu_s"] - b["cpu_s"] for a, b in zip(after["workers"], before["workers"])]
    conns = [a["connections"] - b["connections"] for a, b in zip(after["workers"], before["workers"])]
    return {
        "workers": workers, "mode": mode, "rps": rep["rps"], "p50_us": rep["latency_us"]["p50"], "p99_us": rep["latency_us"]["p99"],
        "worker_spread": round(max(req) / max(1, min(req)), 3), "requests_per_worker": req,
        "connections_per_worker": conns, "cpu_s_per_worker": [round(c, 2) for c in cpu],
    }


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", default="4,8,16")
    ap.add_argument("--modes", default="reuseport,rr,balanced")
    ap.add_argument("--seconds", type=int, default=8)
    ap.add_argument("--repeat", type=int, default=3)
    ap.add_argument("--connections", type=int, default=64)
    ap.add_argument("--path", default="/api/hola")
    ap.add_argument("--out", default="")
    ap.add_argument("extra", nargs="*", help="extra arguments for hola (after --)")
    a = ap.parse_args()
    build()
    cpus = available_cpus()
    rows = []
    for w in [int(x) for x in a.workers.split(",")]:
        if 2 * w > len(cpus):
            print("skip %d workers: needs %d CPUs, have %d" % (w, 2 * w, len(cpus)), file=sys.stderr)
            continue
        server_cpus, load_cpus = cpus[:w], cpus[w:2 * w]
        for mode in a.modes.split(","):
            runs = [one_run(w, mode, a.seconds, server_cpus, load_cpus, a.connections, a.path, a.extra) for _ in range(a.repeat)]
            runs.sort(key=lambda r: r["rps"])
            row = dict(runs[len(runs) // 2])  # the median run, whole
            row["rps_all"] = [round(r["rps"]) for r in runs]
            row["spread_all"] = [r["worker_spread"] for r in runs]
            rows.append(row)
            print("%2d workers %-9s  %9.0f req/s  p50 %4d us  p99 %5d us  spread %.2f  conns %s" %
                  (w, mode, row["rps"], row["p50_us"], row["p99_us"], row["worker_spread"], row["connections_per_worker"]), flush=True)
    doc = {"cpus": len(cpus), "seconds": a.seconds, "repeat": a.repeat, "connections": a.connections, "path": a.path, "rows": rows}
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps(doc, indent=1) + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
