#!/usr/bin/env python
"""GET /api/hola-chaining, every outcome, measured.

  fallback   downstream dead: after the first failures the circuit is open and every request is
             short-circuited to the static fallback (HolaResource.java:112) - compared with
             GET /api/hola on the same cores, which it should now match within 10 %;
  success    downstream is a second `hola --serve-aloha` on its own cores: once with the
             reference's bulkhead of 10 (hystrix.threadpool.default.coreSize, Hystrix default) and
             10 client connections - more would be rejected to the fallback, by design - and once
             with the bulkhead widened to 64 and the usual 64 connections.

Prints requests/sec, latency and the server's CPU per request for each.

    python tools/chaining_paths.py [--seconds 8] [--out FILE]
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

ENV = {"HOSTNAME": None, "ZIPKIN_SERVER_URL": None}


def measure(label, srv, path, connections, load_cpus, seconds, expect):
    before = srv.stats()
    rep = loadgen.run(srv.base_url + path, connections=connections, duration=seconds, warmup=2, threads=min(connections, len(load_cpus)),
                      cpus=load_cpus, expect_body=expect)
    after = srv.stats()
    n = max(1, after["requests"] - before["requests"])
    row = {"case": label, "connections": connections, "rps": rep["rps"], "p50_us": rep["latency_us"]["p50"], "p99_us": rep["latency_us"]["p99"],
           "user_us_per_req": round((after["cpu_user_s"] - before["cpu_user_s"]) * 1e6 / n, 3),
           "sys_us_per_req": round((after["cpu_sys_s"] - before["cpu_sys_s"]) * 1e6 / n, 3),
           "worker_spread": round(max(a["requests"] - b["requests"] for a, b in zip(after["workers"], before["workers"])) /
                                  max(1, min(a["requests"] - b["requests"] for a, b in zip(after["workers"], before["workers"]))), 3)}
    print("%-44s %3d conns %9.0f req/s  p50 %6.1f  p99 %7.1f us  cpu %.2f+%.2f us/req  spread %.2f" %
          (label, connections, row["rps"], row["p50_us"], row["p99_us"], row["user_us_per_req"], row["sys_us_per_req"], row["worker_spread"]), flush=True)
    return row


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=int, default=8)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    build()
    cpus = available_cpus()
    os.sched_setaffinity(0, {cpus[-1]})
    usable = cpus[:-1]
    rows = []
    fallback = b'["Hola de unknown","Aloha response (fallback)"]'

    # 1. Fallback path against /api/hola: server on one half, generator on the other.
    half = len(usable) // 2
    server_cpus, load_cpus = usable[:half], usable[half:]
    cpu_arg = ["--cpus", ",".join(map(str, server_cpus))]
    with HolaServer(args=cpu_arg + ["--aloha-url", "http://127.0.0.1:9/"], env=ENV, threads=len(server_cpus)) as srv:
        rows.append(measure("/api/hola", srv, "/api/hola", 64, load_cpus, a.seconds, b"Hola de unknown"))
        rows.append(measure("/api/hola-chaining, downstream dead (fallback)", srv, "/api/hola-chaining", 64, load_cpus, a.seconds, fallback))
        rows.append(measure("/api/hola (again)", srv, "/api/hola", 64, load_cpus, a.seconds, b"Hola de unknown"))

    # 2. Success path: a third of the CPUs each for hola, the aloha stand-in and the generator.
    third = len(usable) // 3
    server_cpus, aloha_cpus, load_cpus = usable[:third], usable[third:2 * third], usable[2 * third:]
    with HolaServer(args=["--cpus", ",".join(map(str, aloha_cpus)), "--serve-aloha"], env={"HOSTNAME": "next", "ZIPKIN_SERVER_URL": None},
                    threads=len(aloha_cpus)) as aloha:
        expect = b'["Hola de unknown","Aloha mai next"]'
        cpu_arg = ["--cpus", ",".join(map(str, server_cpus)), "--aloha-url", aloha.base_url + "/"]
        with HolaServer(args=cpu_arg, env=ENV, threads=len(server_cpus)) as srv:
            rows.append(measure("success, bulkhead 10 (reference)", srv, "/api/hola-chaining", 10, load_cpus, a.seconds, expect))
        with HolaServer(args=cpu_arg + ["-Dhystrix.threadpool.default.coreSize=64"], env=ENV, threads=len(server_cpus)) as srv:
            rows.append(measure("success, bulkhead widened to 64", srv, "/api/hola-chaining", 64, load_cpus, a.seconds, expect))
            rows.append(measure("/api/hola on the same %d cores" % len(server_cpus), srv, "/api/hola", 64, load_cpus, a.seconds, b"Hola de unknown"))
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({"rows": rows}, indent=1) + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
