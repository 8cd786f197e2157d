"""Drive ``build/hola-load`` and return its JSON report (src/tools/hola_load.cc)."""
from __future__ import annotations

import json
import subprocess
from typing import Dict, Iterable, List, Optional, Sequence

from . import build, paths


def cold_start(command: Sequence[str], port: int, runs: int = 5, path: str = "/api/hola", timeout_ms: int = 30000,
               env: Optional[Dict[str, str]] = None) -> Dict:
    """Times exec() -> first HTTP 200 with ``build/hola-coldstart`` (BASELINE.json config 3).

    ``command`` is the server's full command line; it must listen on ``port``.
    """
    exe = paths().hola_load.parent / "hola-coldstart"
    if not exe.exists():
        build(force=True)
    cmd = [str(exe), "--port", str(port), "--path", path, "--runs", str(runs), "--timeout-ms", str(timeout_ms), "--", *command]
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env,
                          timeout=runs * (timeout_ms / 1000.0 + 5))
    if proc.returncode != 0:
        raise RuntimeError("hola-coldstart failed (%d): %s" % (proc.returncode, proc.stderr[-2000:]))
    return json.loads(proc.stdout.strip().splitlines()[-1])


def run(url: str, connections: int = 64, duration: float = 10.0, warmup: float = 1.0, interval: float = 1.0,
        threads: Optional[int] = None, cpus: Optional[Sequence[int]] = None, rate: Optional[float] = None,
        pipeline: int = 1, expect_status: int = 200, expect_body: Optional[bytes] = None,
        headers: Iterable[str] = (), http10: bool = False, timeout_ms: int = 5000,
        allow_errors: bool = False) -> Dict:
    """Runs one load test and returns the parsed report.

    Raises RuntimeError when the generator saw any error (wrong status, wrong body, broken
    connection, timeout) unless ``allow_errors`` is set: a throughput number from a run with
    failed requests is not a number worth keeping.
    """
    exe = paths().hola_load
    if not exe.exists():
        build()
    cmd: List[str] = [str(exe), "--json", "-c", str(connections), "-d", str(duration), "-w", str(warmup), "-i", str(interval),
                      "--pipeline", str(pipeline), "--expect-status", str(expect_status), "--timeout-ms", str(timeout_ms)]
    if threads:
        cmd += ["-t", str(threads)]
    if cpus:
        cmd += ["--cpus", ",".join(str(c) for c in cpus)]
    if rate:
        cmd += ["--rate", str(rate)]
    if http10:
        cmd += ["--http10"]
    for h in headers:
        cmd += ["-H", h]
    body_file = None
    if expect_body is not None:
        import tempfile
        body_file = tempfile.NamedTemporaryFile(prefix="hola_expect_", delete=False)
        body_file.write(expect_body)
        body_file.close()
        cmd += ["--expect-body-file", body_file.name]
    cmd.append(url)
    try:
        proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                              timeout=duration + warmup + 60)
    finally:
        if body_file is not None:
            import os
            os.unlink(body_file.name)
    if proc.returncode not in (0, 3) or not proc.stdout.strip():
        raise RuntimeError("hola-load failed (%d): %s" % (proc.returncode, proc.stderr[-2000:]))
    report = json.loads(proc.stdout.strip().splitlines()[-1])
    if report["errors"]["total"] and not allow_errors:
        raise RuntimeError("hola-load saw errors: %r" % (report["errors"],))
    return report
