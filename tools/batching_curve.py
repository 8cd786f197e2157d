#!/usr/bin/env python
"""Requests/sec per server core as a function of connections per worker.

Explains the one figure in the scaling record that is not flat: eight small instances (32-64
connections per server thread) serve ~10 % more per core than one big instance (4 connections per
thread) - each system call then carries more requests. Same pinning as bench.py; closed loop.

    python tools/batching_curve.py [--connections 8,16,32,64,128,256,512] [--seconds 5] [--out FILE]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

from hola_b200 import build, loadgen  # noqa: E402
from hola_b200.server import HolaServer  # noqa: E402
from hola_b200.utils import available_cpus  # noqa: E402


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--connections", default="8,16,32,64,128,256,512")
    ap.add_argument("--seconds", type=int, default=5)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    build()
    cpus = available_cpus()
    usable = cpus[:-1] if len(cpus) >= 4 else cpus
    os.sched_setaffinity(0, {cpus[-1]})
    server_cpus, load_cpus = usable[:len(usable) // 2], usable[len(usable) // 2:]
    rows = []
    for backend in ("uring", "epoll"):
        for c in [int(x) for x in a.connections.split(",")]:
            with HolaServer(args=["--cpus", ",".join(map(str, server_cpus)), "--io-backend", backend],
                            env={"HOSTNAME": None, "ZIPKIN_SERVER_URL": None}, threads=len(server_cpus)) as srv:
                before = srv.stats()
                rep = loadgen.run(srv.base_url + "/api/hola", connections=c, duration=a.seconds, warmup=1,
                                  threads=min(c, len(load_cpus)), cpus=load_cpus, expect_body=b"Hola de unknown")
                after = srv.stats()
            n = max(1, after["requests"] - before["requests"])
            row = {"backend": after["io_backend"], "connections": c, "per_worker": round(c / len(server_cpus), 1), "rps": rep["rps"],
                   "per_core_rps": round(rep["rps"] / len(server_cpus)), "p50_us": rep["latency_us"]["p50"], "p99_us": rep["latency_us"]["p99"],
                   "user_us": round((after["cpu_user_s"] - before["cpu_user_s"]) * 1e6 / n, 3),
                   "sys_us": round((after["cpu_sys_s"] - before["cpu_sys_s"]) * 1e6 / n, 3)}
            rows.append(row)
            print("%-5s %4d connections (%5.1f per worker) %9.0f req/s  %6d per core  p50 %6.1f  p99 %6.1f us  cpu %.2f+%.2f us/req" %
                  (row["backend"], c, row["per_worker"], row["rps"], row["per_core_rps"], row["p50_us"], row["p99_us"], row["user_us"], row["sys_us"]), flush=True)
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({"server_cpus": server_cpus, "load_cpus": load_cpus, "rows": rows}, indent=1) + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
