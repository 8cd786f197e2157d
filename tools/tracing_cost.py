#!/usr/bin/env python
"""What tracing costs when spans are really reported.

Round 1 measured tracing with a collector nobody runs: every span took the ring-full branch, so
"tracing adds 0.01 us" was the price of a failed push. Here a live collector takes the spans
(`hola-null` answers every POST with a 200 - a collector that is never the bottleneck), and the
service is driven

  * open loop at a series of request rates, to find the highest rate at which no span is dropped
    and what a *reported* span costs in CPU (server user+sys per request, tracing on vs off);
  * closed loop at saturation, to see the drop rate there,

once with the reference's queue sizes (1024-span rings, 1000 spans per POST, 1 s flush:
ZipkinResource.java:33-41 builds Brave's HttpSpanCollector with its defaults) and once with large
rings (--zipkin-ring 65536 --zipkin-max-batch 10000 --zipkin-flush-ms 100).

    python tools/tracing_cost.py [--seconds 5] [--out FILE]
"""
from __future__ import annotations

import argparse
import json
import os
import signal
import subprocess
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

from hola_b200 import build, loadgen, paths  # noqa: E402
from hola_b200.reference import free_port  # noqa: E402
from hola_b200.server import HolaServer  # noqa: E402
from hola_b200.utils import available_cpus  # noqa: E402


def run_case(label, server_cpus, load_cpus, seconds, zipkin_url, extra, rate):
    env = {"HOSTNAME": None, "ZIPKIN_SERVER_URL": zipkin_url}
    with HolaServer(args=["--cpus", ",".join(map(str, server_cpus))] + extra, env=env, threads=len(server_cpus)) as srv:
        before = srv.stats()
        rep = loadgen.run(srv.base_url + "/api/hola", connections=64, duration=seconds, warmup=1, threads=min(64, len(load_cpus)),
                          cpus=load_cpus, expect_body=b"Hola de unknown", rate=rate)
        time.sleep(1.5)  # let the reporter finish what is queued
        after = srv.stats()
    n = max(1, after["requests"] - before["requests"])
    spans = {k: after["spans"][k] - before["spans"][k] for k in ("reported", "dropped", "posts_ok", "posts_failed")}
    row = {"case": label, "offered_rps": rate, "rps": rep["rps"], "p50_us": rep["latency_us"]["p50"], "p99_us": rep["latency_us"]["p99"],
           "requests": n, "spans": spans, "drop_pct": round(100.0 * spans["dropped"] / max(1, spans["dropped"] + spans["reported"]), 2),
           "user_us_per_req": round((after["cpu_user_s"] - before["cpu_user_s"]) * 1e6 / n, 3),
           "sys_us_per_req": round((after["cpu_sys_s"] - before["cpu_sys_s"]) * 1e6 / n, 3)}
    print("%-28s offered %8s  got %9.0f req/s  p99 %6.1f us  spans %9d reported %9d dropped (%5.1f %%)  cpu %.2f+%.2f us/req" %
          (label, rate or "max", row["rps"], row["p99_us"], spans["reported"], spans["dropped"], row["drop_pct"],
           row["user_us_per_req"], row["sys_us_per_req"]), flush=True)
    return row


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=int, default=5)
    ap.add_argument("--rates", default="25000,50000,100000,200000,400000,800000")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    build()
    cpus = available_cpus()
    # last CPU: this script; the one before: the collector; the rest: server half, generator half.
    os.sched_setaffinity(0, {cpus[-1]})
    collector_cpu = cpus[-2]
    usable = cpus[:-2]
    server_cpus, load_cpus = usable[:len(usable) // 2], usable[len(usable) // 2:]
    port = free_port()
    collector = subprocess.Popen([str(paths().hola.parent / "hola-null"), "--port", str(port), "--threads", "1", "--cpus", str(collector_cpu)],
                                 stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    time.sleep(0.3)
    url = "http://127.0.0.1:%d" % port
    big = ["--zipkin-ring", "65536", "--zipkin-max-batch", "10000", "--zipkin-flush-ms", "100"]
    rows = []
    try:
        for rate in [int(x) for x in a.rates.split(",")]:
            rows.append(run_case("tracing off", server_cpus, load_cpus, a.seconds, None, [], rate))
            rows.append(run_case("reference queue sizes", server_cpus, load_cpus, a.seconds, url, [], rate))
            rows.append(run_case("large rings", server_cpus, load_cpus, a.seconds, url, big, rate))
        rows.append(run_case("tracing off", server_cpus, load_cpus, a.seconds, None, [], None))
        rows.append(run_case("reference queue sizes", server_cpus, load_cpus, a.seconds, url, [], None))
        rows.append(run_case("large rings", server_cpus, load_cpus, a.seconds, url, big, None))
        rows.append(run_case("dead collector (as-is bench)", server_cpus, load_cpus, a.seconds, "http://127.0.0.1:9", [], None))
    finally:
        collector.send_signal(signal.SIGTERM)
        collector.wait(10)
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({"server_cpus": server_cpus, "load_cpus": load_cpus, "collector_cpu": collector_cpu, "rows": rows}, indent=1) + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
