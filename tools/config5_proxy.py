#!/usr/bin/env python
"""BASELINE.json config 5, as close as a box without a container runtime allows.

Config 5 is "container image built from the Dockerfile, same 64-client load test inside the
container". No machine this project sees has docker/podman, so the image (Dockerfile: FROM scratch,
/app/hola, EXPOSE 8080, ENV ZIPKIN_SERVER_URL=...; the reference's is /root/reference/Dockerfile:1-10)
was never built. What the image *contains* can be run, though: the static executable alone in an
empty directory entered with chroot(2) - no /etc, /proc, /sys, libraries or resolver files - with
the environment the image sets, under the load of config 4 with bench.py's CPU plan.

It is a proxy, not a container: no namespaces, no cgroup limits, no overlay filesystem, the host's
network stack without a veth pair in between. Needs root (CAP_SYS_CHROOT).

    python tools/config5_proxy.py [--seconds 60] [--out FILE]
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import signal
import subprocess
import sys
import tempfile
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

import hola_b200  # noqa: E402
from hola_b200 import client, loadgen  # noqa: E402
from hola_b200.reference import free_port  # noqa: E402
from hola_b200.utils import available_cpus  # noqa: E402


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=int, default=60)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    static = hola_b200.build("static").hola
    cpus = available_cpus()
    usable = cpus[:-1] if len(cpus) >= 4 else cpus
    server_cpus, load_cpus = usable[:len(usable) // 2], usable[len(usable) // 2:]
    os.sched_setaffinity(0, {cpus[-1]})
    root = Path(tempfile.mkdtemp(prefix="hola_rootfs_"))
    try:
        (root / "app").mkdir()
        shutil.copy2(static, root / "app" / "hola")
        contents = sorted(str(p.relative_to(root)) for p in root.rglob("*"))
        port = free_port()
        env = {"ZIPKIN_SERVER_URL": "http://zipkin-query:9411"}  # Dockerfile: ENV ZIPKIN_SERVER_URL; nothing else is set
        proc = subprocess.Popen(["/app/hola", "--port", str(port), "--bind", "127.0.0.1", "--threads", str(len(server_cpus)),
                                 "--cpus", ",".join(map(str, server_cpus))], env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True,
                                preexec_fn=lambda: (os.chroot(str(root)), os.chdir("/")))
        try:
            t0 = time.monotonic()
            while True:
                try:
                    if client.get("127.0.0.1", port, "/api/health", timeout=1).status == 200:
                        break
                except OSError:
                    if time.monotonic() - t0 > 5:
                        raise
                    time.sleep(0.001)
            chained = client.get("127.0.0.1", port, "/api/hola-chaining").body.decode()
            rep = loadgen.run("http://127.0.0.1:%d/api/hola" % port, connections=64, duration=a.seconds, warmup=3, threads=min(64, len(load_cpus)),
                              cpus=load_cpus, expect_body=b"Hola de unknown")
        finally:
            proc.send_signal(signal.SIGTERM)
            _, err = proc.communicate(timeout=15)
        out = {"what": "config-5 proxy: static executable alone in an empty chroot (not a container)", "rootfs": contents,
               "exit_code": proc.returncode, "startup_line": err.splitlines()[0] if err else None,
               "chaining_without_any_resolver_file": chained, "rps": rep["rps"], "latency_us": rep["latency_us"], "requests": rep["requests"],
               "errors": rep["errors"]["total"], "seconds": a.seconds, "server_cpus": server_cpus, "load_cpus": load_cpus}
    finally:
        shutil.rmtree(root, ignore_errors=True)
    print(json.dumps(out))
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps(out, indent=1) + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
