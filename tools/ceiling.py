#!/usr/bin/env python
"""The TCP ceiling of this machine, measured.

Runs the identical hola-load configuration (64 keep-alive connections, closed loop, bodies
verified, same pinning as bench.py) against

  * `hola --io-backend epoll`     the service on its readiness backend (tracing off and as-is),
  * `hola --io-backend uring`     the service on its io_uring backend,
  * `hola-null --io-backend epoll`  recv -> canned 200 -> send with hola's threading: no parsing at all,
  * `hola-null --io-backend uring`  the same probe on io_uring (multishot recv, batched sends),

alternating between them `--repeat` times, and prints requests/sec, latency and the server's CPU
time per request (user/sys, from /proc/<pid>/stat). `hola / null-epoll` is the fraction of the
machine's ceiling the service reaches with its system-call pattern; `null-uring / null-epoll` is
the most an io_uring backend could add. The reference's I/O layer (Undertow/XNIO,
/root/reference/pom.xml:122-129) uses the epoll pattern.

    python tools/ceiling.py [--seconds 8] [--repeat 3] [--workers N] [--out FILE]
"""
from __future__ import annotations

import argparse
import json
import os
import signal
import statistics
import subprocess
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

from hola_b200 import build, loadgen, paths  # noqa: E402
from hola_b200.reference import free_port  # noqa: E402
from hola_b200.utils import available_cpus  # noqa: E402

TICK = os.sysconf("SC_CLK_TCK")


def cpu_times(pid: int):
    f = open("/proc/%d/stat" % pid).read().rsplit(")", 1)[1].split()
    return int(f[11]) / TICK, int(f[12]) / TICK


def wait_listening(port: int, timeout: float = 5.0) -> None:
    import socket
    deadline = time.monotonic() + timeout
    while time.monotonic() < deadline:
        try:
            socket.create_connection(("127.0.0.1", port), timeout=0.5).close()
            return
        except OSError:
            time.sleep(0.01)
    raise RuntimeError("nothing listens on %d" % port)


def one(label: str, cmd, env, port: int, seconds: int, load_cpus, path: str):
    proc = subprocess.Popen(cmd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    try:
        wait_listening(port)
        u0, s0 = cpu_times(proc.pid)
        rep = loadgen.run("http://127.0.0.1:%d%s" % (port, path), connections=64, duration=seconds, warmup=1,
                          threads=min(len(load_cpus), 64), cpus=load_cpus, expect_body=b"Hola de unknown")
        u1, s1 = cpu_times(proc.pid)
    finally:
        proc.send_signal(signal.SIGTERM)
        try:
            proc.wait(10)
        except subprocess.TimeoutExpired:
            proc.kill()
    total = rep["requests"] * (seconds + 1) / seconds  # CPU was also spent during the warm-up second
    return {"label": label, "rps": rep["rps"], "p50_us": rep["latency_us"]["p50"], "p99_us": rep["latency_us"]["p99"],
            "user_us_per_req": round((u1 - u0) * 1e6 / total, 3), "sys_us_per_req": round((s1 - s0) * 1e6 / total, 3)}


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=int, default=8)
    ap.add_argument("--repeat", type=int, default=3)
    ap.add_argument("--workers", type=int, default=0, help="server threads (default: as bench.py splits the machine)")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    build()
    cpus = available_cpus()
    usable = cpus[:-1] if len(cpus) >= 4 else cpus  # the last CPU stays free, as in bench.py
    n_server = a.workers or len(usable) // 2
    server_cpus, load_cpus = usable[:n_server], usable[n_server:]
    os.sched_setaffinity(0, {cpus[-1]})
    cpu_arg = ",".join(map(str, server_cpus))
    bin_dir = paths().hola.parent
    base_env = {k: v for k, v in os.environ.items() if k not in ("HOSTNAME", "ZIPKIN_SERVER_URL")}
    variants = [
        ("hola (tracing off)", lambda p: [str(bin_dir / "hola"), "--port", str(p), "--bind", "127.0.0.1", "--threads", str(n_server), "--cpus", cpu_arg, "--quiet",
                                          "--io-backend", "epoll"], base_env),
        ("hola (as-is)", lambda p: [str(bin_dir / "hola"), "--port", str(p), "--bind", "127.0.0.1", "--threads", str(n_server), "--cpus", cpu_arg, "--quiet",
                                    "--io-backend", "epoll"], dict(base_env, ZIPKIN_SERVER_URL="http://127.0.0.1:9")),
        ("hola uring (tracing off)", lambda p: [str(bin_dir / "hola"), "--port", str(p), "--bind", "127.0.0.1", "--threads", str(n_server), "--cpus", cpu_arg, "--quiet",
                                                "--io-backend", "uring"], base_env),
        ("hola uring (as-is)", lambda p: [str(bin_dir / "hola"), "--port", str(p), "--bind", "127.0.0.1", "--threads", str(n_server), "--cpus", cpu_arg, "--quiet",
                                          "--io-backend", "uring"], dict(base_env, ZIPKIN_SERVER_URL="http://127.0.0.1:9")),
        ("null epoll", lambda p: [str(bin_dir / "hola-null"), "--port", str(p), "--threads", str(n_server), "--cpus", cpu_arg, "--io-backend", "epoll"], base_env),
        ("null uring", lambda p: [str(bin_dir / "hola-null"), "--port", str(p), "--threads", str(n_server), "--cpus", cpu_arg, "--io-backend", "uring"], base_env),
    ]
    runs = {label: [] for label, _, _ in variants}
    for _ in range(a.repeat):
        for label, make_cmd, env in variants:
            port = free_port()
            try:
                r = one(label, make_cmd(port), env, port, a.seconds, load_cpus, "/api/hola")
            except Exception as exc:  # noqa: BLE001
                print("%-26s failed: %s" % (label, exc), flush=True)
                continue
            runs[label].append(r)
            print("%-26s %9.0f req/s  p50 %5.1f us  p99 %6.1f us  user %.2f  sys %.2f us/req" %
                  (label, r["rps"], r["p50_us"], r["p99_us"], r["user_us_per_req"], r["sys_us_per_req"]), flush=True)
    summary = {}
    for label, rs in runs.items():
        if rs:
            summary[label] = {k: round(statistics.median(r[k] for r in rs), 3) for k in ("rps", "p50_us", "p99_us", "user_us_per_req", "sys_us_per_req")}
    ratios = {}
    if "null epoll" in summary:
        for label in ("hola (tracing off)", "hola (as-is)", "hola uring (tracing off)", "hola uring (as-is)", "null uring"):
            if label in summary:
                ratios[label + " / null epoll"] = round(summary[label]["rps"] / summary["null epoll"]["rps"], 3)
    doc = {"server_cpus": server_cpus, "load_cpus": load_cpus, "seconds": a.seconds, "repeat": a.repeat, "median": summary, "ratios": ratios, "runs": runs}
    for a_label, b_label in (("hola uring (tracing off)", "hola (tracing off)"), ("hola uring (as-is)", "hola (as-is)")):
        if a_label in summary and b_label in summary:
            ratios[a_label + " / " + b_label] = round(summary[a_label]["rps"] / summary[b_label]["rps"], 3)
            ratios["p99: " + a_label + " / " + b_label] = round(summary[a_label]["p99_us"] / summary[b_label]["p99_us"], 3)
    print(json.dumps({"median": summary, "ratios": ratios}, indent=1))
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps(doc, indent=1) + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
