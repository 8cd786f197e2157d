"""Launching the *reference* service (the Java uber-jar) for the benchmark's reference arm.

How the reference is run is stated in /root/reference/Readme.md:13-16 (``mvn wildfly-swarm:run`` or
``java -jar target/hola-swarm.jar``) and /root/reference/Dockerfile:3-6 (``JAVA_APP_JAR``,
``JAVA_OPTIONS=-Xmx512m``, ``ZIPKIN_SERVER_URL`` must be set or the deployment fails, SURVEY.md
2.1 C9). No box this project has seen carries a JVM, so nothing here has ever launched the real
jar; the code path is exercised by the tests with a stand-in command that is *not* this
repository's server (a few lines of Python ``http.server``).

Resolution order of the command:

1. ``$HOLA_REFERENCE_CMD`` - a command line; ``{port}`` is replaced by the port to listen on.
2. ``java`` on PATH and ``baseline/_ref/hola-swarm.jar`` present:
   ``java -Xmx512m -Dswarm.http.port={port} -Dswarm.bind.address=127.0.0.1 -jar <jar>``.

Anything else: :func:`reference_command` returns ``(None, why)``.
"""
from __future__ import annotations

import os
import shlex
import shutil
import socket
import subprocess
import time
from pathlib import Path
from typing import Dict, List, Optional, Sequence, Tuple

from . import REPO_ROOT, client

REFERENCE_JAR = REPO_ROOT / "baseline" / "_ref" / "hola-swarm.jar"


def reference_command() -> Tuple[Optional[List[str]], str]:
    """The reference's launch command as an argv template, or ``(None, one-line reason)``."""
    override = os.environ.get("HOLA_REFERENCE_CMD", "").strip()
    if override:
        return shlex.split(override), "HOLA_REFERENCE_CMD"
    java = shutil.which("java")
    if java and REFERENCE_JAR.exists():
        return [java, "-Xmx512m", "-Dswarm.http.port={port}", "-Dswarm.bind.address=127.0.0.1", "-jar", str(REFERENCE_JAR)], "jar"
    missing = [tool for tool in ("java", "mvn", "docker") if shutil.which(tool) is None]
    if java and not REFERENCE_JAR.exists():
        return None, ("a JVM exists but %s does not: building it needs Maven plus the WildFly Swarm 1.0.0.Final artefacts from "
                      "Maven Central (no network); set HOLA_REFERENCE_CMD or drop the jar there" % REFERENCE_JAR)
    return None, ("reference is a Java EE/WildFly Swarm Maven project; no %s on this image, no ~/.m2, no network to fetch Swarm "
                  "1.0.0.Final; pip install of /root/reference fails (not a Python project); set HOLA_REFERENCE_CMD to run one"
                  % "/".join(missing or ["jar"]))


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class ReferenceServer:
    """One reference process on its own port, pinned to ``cpus``; same surface as HolaServer as far
    as the benchmark uses it (``base_url``, ``startup_seconds``, ``stats()`` -> None, ``stop()``)."""

    def __init__(self, command: Sequence[str], cpus: Optional[Sequence[int]] = None,
                 env: Optional[Dict[str, Optional[str]]] = None, ready_path: str = "/api/hola", start_timeout: float = 180.0):
        self.port = free_port()
        self.command = [part.replace("{port}", str(self.port)) for part in command]
        self.cpus = list(cpus or [])
        self.env_overrides = dict(env or {})
        self.ready_path = ready_path
        self.start_timeout = start_timeout
        self.proc: Optional[subprocess.Popen] = None
        self.startup_seconds = 0.0
        self.threads = len(self.cpus) or None
        self._log: Optional[Path] = None
        self.keep_log = False  # True: the log file survives stop(), so that log_tail() still works afterwards

    def start(self) -> "ReferenceServer":
        env = dict(os.environ)
        for key, value in self.env_overrides.items():
            if value is None:
                env.pop(key, None)
            else:
                env[key] = value
        cpus = self.cpus

        def pin() -> None:
            os.setsid()
            if cpus:
                os.sched_setaffinity(0, cpus)

        import tempfile
        log = tempfile.NamedTemporaryFile(prefix="hola_reference_", suffix=".log", delete=False)
        self._log = Path(log.name)
        t0 = time.monotonic()
        self.proc = subprocess.Popen(self.command, env=env, stdout=log, stderr=subprocess.STDOUT, preexec_fn=pin)
        log.close()
        deadline = t0 + self.start_timeout
        while time.monotonic() < deadline:
            if self.proc.poll() is not None:
                raise RuntimeError("reference exited with %s before answering: %s" % (self.proc.returncode, self.log_tail()))
            try:
                r = client.get("127.0.0.1", self.port, self.ready_path, timeout=2.0)
                if r.status == 200:
                    self.startup_seconds = time.monotonic() - t0
                    return self
            except OSError:
                pass
            time.sleep(0.02)
        self.stop()
        raise RuntimeError("reference did not answer %s with 200 within %.0f s: %s" % (self.ready_path, self.start_timeout, self.log_tail()))

    def log_tail(self, n: int = 2000) -> str:
        try:
            return self._log.read_text(errors="replace")[-n:] if self._log else ""
        except OSError:
            return ""

    def stats(self):
        return None  # the reference has no counter dump

    def stop(self, timeout: float = 20.0) -> int:
        if self.proc is None:
            return 0
        if self.proc.poll() is None:
            try:
                os.killpg(self.proc.pid, 15)  # the whole group: `mvn wildfly-swarm:run` forks a JVM
            except ProcessLookupError:
                pass
            try:
                self.proc.wait(timeout)
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(self.proc.pid, 9)
                except ProcessLookupError:
                    pass
                self.proc.wait()
        code = self.proc.returncode
        self.proc = None
        if self._log and not self.keep_log:
            try:
                self._log.unlink()
            except OSError:
                pass
        return code

    def __enter__(self) -> "ReferenceServer":
        return self.start()

    def __exit__(self, *exc) -> None:
        self.stop()

    @property
    def base_url(self) -> str:
        return "http://127.0.0.1:%d" % self.port
