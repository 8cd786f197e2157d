"""Run the ``hola`` executable as a child process (what ``java -jar hola-swarm.jar`` is to the
reference, /root/reference/Readme.md:13-16)."""
from __future__ import annotations

import os
import re
import signal
import subprocess
import threading
import time
from typing import Dict, Iterable, List, Optional

from . import build, paths

_LISTEN_RE = re.compile(r"listening on (\S+):(\d+) with (\d+) thread")


class HolaServer:
    """Context manager around one ``build/hola`` process.

    ``env`` *replaces* the variables it names; a value of ``None`` removes the variable (the
    reference's greeting depends on whether ``HOSTNAME`` is exported at all, HolaResource.java:57).
    """

    def __init__(self, args: Iterable[str] = (), env: Optional[Dict[str, Optional[str]]] = None, port: int = 0,
                 threads: Optional[int] = 2, variant: str = "release", start_timeout: float = 10.0):
        self.args: List[str] = list(args)
        self.env_overrides = dict(env or {})
        self.requested_port = port
        self.threads = threads
        self.variant = variant
        self.start_timeout = start_timeout
        self.proc: Optional[subprocess.Popen] = None
        self.port: int = 0
        self.startup_seconds: float = 0.0
        self.stderr_lines: List[str] = []
        self.stdout_lines: List[str] = []
        self._readers: List[threading.Thread] = []
        self._listening = threading.Event()

    # -- lifecycle ---------------------------------------------------------------------------
    def start(self) -> "HolaServer":
        exe = paths(self.variant).hola
        if not exe.exists():
            build(self.variant)
        env = dict(os.environ)
        for key, value in self.env_overrides.items():
            if value is None:
                env.pop(key, None)
            else:
                env[key] = value
        cmd = [str(exe), "--port", str(self.requested_port), "--bind", "127.0.0.1"]
        if self.threads:
            cmd += ["--threads", str(self.threads)]
        cmd += self.args
        t0 = time.monotonic()
        self.proc = subprocess.Popen(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                                     encoding="utf-8", errors="replace")
        self._readers = [
            threading.Thread(target=self._pump, args=(self.proc.stderr, self.stderr_lines, True), daemon=True),
            threading.Thread(target=self._pump, args=(self.proc.stdout, self.stdout_lines, False), daemon=True),
        ]
        for r in self._readers:
            r.start()
        if not self._listening.wait(self.start_timeout):
            self.stop()
            raise RuntimeError("hola did not start: " + "".join(self.stderr_lines)[-2000:])
        if self.port == 0:
            # stderr reached EOF without a "listening" line: the process is on its way out.
            try:
                self.proc.wait(5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
                self.proc.wait()
            code = self.proc.returncode
            self.proc = None
            raise RuntimeError("hola exited with %s before listening: %s" % (code, "".join(self.stderr_lines)[-2000:]))
        self.startup_seconds = time.monotonic() - t0
        return self

    def _pump(self, stream, sink: List[str], watch: bool) -> None:
        for line in stream:
            sink.append(line)
            if watch and not self._listening.is_set():
                m = _LISTEN_RE.search(line)
                if m:
                    self.port = int(m.group(2))
                    self._listening.set()
        if watch:
            self._listening.set()  # process ended; unblock start()

    def stop(self, timeout: float = 10.0) -> int:
        if self.proc is None:
            return 0
        if self.proc.poll() is None:
            self.proc.send_signal(signal.SIGTERM)
            try:
                self.proc.wait(timeout)
            except subprocess.TimeoutExpired:
                self.proc.kill()
                self.proc.wait()
        for r in self._readers:
            r.join(timeout=2)
        code = self.proc.returncode
        self.proc = None
        return code

    def __enter__(self) -> "HolaServer":
        return self.start()

    def __exit__(self, *exc) -> None:
        self.stop()

    def stats(self, timeout: float = 5.0) -> Dict:
        """Asks the running process for its counters (SIGUSR1 -> one ``hola-stats {json}`` line on
        stderr): totals, CPU seconds, span counts and one entry per worker thread."""
        import json
        if self.proc is None or self.proc.poll() is not None:
            raise RuntimeError("hola is not running")
        seen = sum(1 for line in self.stderr_lines if line.startswith("hola-stats "))
        self.proc.send_signal(signal.SIGUSR1)
        deadline = time.monotonic() + timeout
        while time.monotonic() < deadline:
            lines = [line for line in self.stderr_lines if line.startswith("hola-stats ")]
            if len(lines) > seen:
                return json.loads(lines[-1][len("hola-stats "):])
            time.sleep(0.005)
        raise RuntimeError("no hola-stats line within %.1f s" % timeout)

    # -- conveniences ------------------------------------------------------------------------
    @property
    def base_url(self) -> str:
        return "http://127.0.0.1:%d" % self.port

    @property
    def stderr(self) -> str:
        return "".join(self.stderr_lines)

    @property
    def stdout(self) -> str:
        return "".join(self.stdout_lines)
