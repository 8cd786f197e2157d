"""Shared fixtures: native build, server launcher, fake downstream services.

The reference has no tests at all, so the oracle is the observable contract
of SURVEY.md 2.2, checked black-box against the running binary, plus unit tests of the building
blocks through the C ABI.

Markers: ``gpu`` is registered because the driver selects on it, but nothing here carries it -
no test needs a GPU, since the project contains no GPU code. ``pytest -m gpu``
therefore selects nothing; that empty selection is reported as success (see
``pytest_sessionfinish``) rather than as pytest's "no tests collected" error code.
"""
from __future__ import annotations

import json
import os
import sys
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from pathlib import Path
from typing import Dict, List, Optional

import pytest

REPO = Path(__file__).resolve().parent.parent
if str(REPO) not in sys.path:
    sys.path.insert(0, str(REPO))

import hola_b200  # noqa: E402
from hola_b200.server import HolaServer  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (none of this project's tests do: it has no GPU code)")
    config.addinivalue_line("markers", "slow: takes more than a few seconds")


def pytest_report_header(config):
    """Build the native artefacts and load libhola.so before anything is collected.

    Every test exercises native code (the executable or the library); if it cannot be built or
    loaded the session should say so at the top instead of failing test by test.
    """
    from hola_b200 import native
    hola_b200.build()
    return "hola native library: %s (version %s)" % (hola_b200.paths().libhola, native.version())


def pytest_addoption(parser):
    parser.addoption("--variant", default="release", choices=["release", "asan", "tsan", "static", "cov"],
                     help="which build of the hola executable the black-box tests launch "
                          "(asan: AddressSanitizer+UBSan, tsan: ThreadSanitizer; SURVEY.md 5.2)")


def pytest_sessionfinish(session, exitstatus):
    # `-m gpu` deselects everything by design; do not let that read as a failure.
    markexpr = getattr(session.config.option, "markexpr", "") or ""
    if exitstatus == pytest.ExitCode.NO_TESTS_COLLECTED and markexpr.strip() == "gpu":
        tr = session.config.pluginmanager.get_plugin("terminalreporter")
        if tr is not None:
            tr.write_line("no GPU tests exist: the project is CPU-only (SURVEY.md 2.3); an empty `-m gpu` selection is expected")
        session.exitstatus = pytest.ExitCode.OK


@pytest.fixture(scope="session", autouse=True)
def native_build():
    """Everything is tested against the real native artefacts; build them once per session."""
    return hola_b200.build()


@pytest.fixture
def hola_server(request):
    """Factory: ``hola_server(args=[...], env={...}, threads=2)`` -> started HolaServer.

    With ``--variant asan|tsan`` the sanitizer build is launched instead, and the teardown fails
    the test if the sanitizer reported anything or the process did not exit cleanly.
    """
    variant = request.config.getoption("--variant")
    if variant != "release":
        hola_b200.build(variant)
    started: List[HolaServer] = []

    def start(args=(), env: Optional[Dict[str, Optional[str]]] = None, threads: int = 2, **kw) -> HolaServer:
        merged: Dict[str, Optional[str]] = {"HOSTNAME": None, "ZIPKIN_SERVER_URL": None, "hello": None}
        if variant == "asan":
            merged["ASAN_OPTIONS"] = "detect_leaks=1:abort_on_error=0:exitcode=66"
            merged["UBSAN_OPTIONS"] = "print_stacktrace=1:halt_on_error=1:exitcode=66"
        elif variant == "tsan":
            merged["TSAN_OPTIONS"] = "halt_on_error=0:exitcode=66:second_deadlock_stack=1"
        merged.update(env or {})
        kw.setdefault("start_timeout", 10.0 if variant == "release" else 30.0)
        srv = HolaServer(args=args, env=merged, threads=threads, variant=variant, **kw).start()
        started.append(srv)
        return srv

    yield start
    problems = []
    for srv in started:
        code = srv.stop()
        if variant in ("asan", "tsan"):
            if code not in (0, None) or "Sanitizer" in srv.stderr or "runtime error:" in srv.stderr:
                problems.append("exit %s\n%s" % (code, srv.stderr[-6000:]))
    assert not problems, "sanitizer findings:\n" + "\n----\n".join(problems)


class _FakeServer(ThreadingHTTPServer):
    # The default accept backlog of 5 drops simultaneous connects when the machine is busy.
    request_queue_size = 128
    daemon_threads = True


class _Recorder:
    def __init__(self):
        self.lock = threading.Lock()
        self.requests: List[Dict] = []

    def add(self, item: Dict) -> None:
        with self.lock:
            self.requests.append(item)

    def snapshot(self) -> List[Dict]:
        with self.lock:
            return list(self.requests)


class FakeAloha(_Recorder):
    """Stand-in for the downstream `aloha` service (GET /api/aloha-chaining -> JSON list)."""

    def __init__(self):
        super().__init__()
        self.mode = "ok"
        self.payload = ["Aloha mai fake-aloha", "Olá de fake-ola", "Bonjour de fake-bonjour"]
        self.delay = 0.0
        self.fail_first = 0  # drop this many upcoming requests without answering
        self.raw = []        # mode "raw": byte strings written one after the other
        fake = self

        class Handler(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"

            def log_message(self, *a):
                pass

            def do_GET(self):
                fake.add({"path": self.path, "headers": {k.lower(): v for k, v in self.headers.items()}, "time": time.time()})
                if fake.delay:
                    time.sleep(fake.delay)
                mode = fake.mode
                with fake.lock:
                    if fake.fail_first > 0:
                        fake.fail_first -= 1
                        mode = "close"
                if mode == "ok":
                    body = json.dumps(fake.payload).encode()
                    self._reply(200, body, "application/json")
                elif mode == "chunked":
                    body = json.dumps(fake.payload).encode()
                    self.send_response(200)
                    self.send_header("Content-Type", "application/json")
                    self.send_header("Transfer-Encoding", "chunked")
                    self.end_headers()
                    half = len(body) // 2
                    for part in (body[:half], body[half:]):
                        self.wfile.write(b"%x\r\n" % len(part) + part + b"\r\n")
                    self.wfile.write(b"0\r\n\r\n")
                elif mode == "500":
                    self._reply(500, b"boom", "text/plain")
                elif mode == "404":
                    self._reply(404, b"nope", "text/plain")
                elif mode == "garbage":
                    self._reply(200, b"<html>not json</html>", "text/html")
                elif mode == "object":
                    self._reply(200, b'{"greeting":"aloha"}', "application/json")
                elif mode == "close":
                    self.close_connection = True
                    self.connection.close()
                elif mode == "raw":
                    # Arbitrary bytes (possibly in several writes), then the connection closes.
                    for part in fake.raw:
                        self.wfile.write(part)
                        self.wfile.flush()
                        time.sleep(0.02)
                    self.close_connection = True
                else:
                    self._reply(200, b"[]", "application/json")

            def _reply(self, status, body, ctype):
                self.send_response(status)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

        self.httpd = _FakeServer(("127.0.0.1", 0), Handler)
        self.httpd.daemon_threads = True
        self.port = self.httpd.server_address[1]
        self.thread = threading.Thread(target=self.httpd.serve_forever, daemon=True)
        self.thread.start()

    @property
    def url(self) -> str:
        return "http://127.0.0.1:%d/" % self.port

    def stop(self):
        self.httpd.shutdown()
        self.httpd.server_close()


class FakeZipkin(_Recorder):
    """Stand-in for the Zipkin collector: records what is POSTed to /api/v1/spans."""

    def __init__(self):
        super().__init__()
        self.status = 202
        fake = self

        class Handler(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"

            def log_message(self, *a):
                pass

            def do_POST(self):
                n = int(self.headers.get("Content-Length", "0"))
                body = self.rfile.read(n)
                fake.add({"path": self.path, "content_type": self.headers.get("Content-Type"), "body": body})
                self.send_response(fake.status)
                self.send_header("Content-Length", "0")
                self.end_headers()

        self.httpd = _FakeServer(("127.0.0.1", 0), Handler)
        self.httpd.daemon_threads = True
        self.port = self.httpd.server_address[1]
        self.thread = threading.Thread(target=self.httpd.serve_forever, daemon=True)
        self.thread.start()

    @property
    def url(self) -> str:
        return "http://127.0.0.1:%d" % self.port

    def spans(self) -> List[Dict]:
        out: List[Dict] = []
        for rq in self.snapshot():
            out.extend(json.loads(rq["body"]))
        return out

    def wait_for_spans(self, count: int, timeout: float = 5.0) -> List[Dict]:
        deadline = time.time() + timeout
        while time.time() < deadline:
            spans = self.spans()
            if len(spans) >= count:
                return spans
            time.sleep(0.05)
        return self.spans()

    def stop(self):
        self.httpd.shutdown()
        self.httpd.server_close()


@pytest.fixture
def fake_aloha():
    f = FakeAloha()
    yield f
    f.stop()


@pytest.fixture
def fake_zipkin():
    f = FakeZipkin()
    yield f
    f.stop()
