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
    parser.addoption("--io-backend", default="auto", choices=["auto", "epoll", "uring"],
                     help="run every launched server with this --io-backend (auto, the service's default: io_uring where the "
                          "kernel offers it, else epoll). test_server_core.py and test_chaining.py always run on both.")


_deselected = {"count": 0}


def pytest_deselected(items):
    _deselected["count"] += len(items)


def pytest_sessionfinish(session, exitstatus):
    # `-m gpu` deselects everything by design; do not let that read as a failure. Only a selection
    # that really was "all tests collected, every one of them deselected" qualifies: a session that
    # collected nothing at all (an import error, a wrong path) keeps pytest's error code.
    markexpr = getattr(session.config.option, "markexpr", "") or ""
    collected_fine = _deselected["count"] >= 100 and not getattr(session, "shouldfail", False)
    if exitstatus == pytest.ExitCode.NO_TESTS_COLLECTED and markexpr.strip() == "gpu" and collected_fine:
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
    io_backend = request.config.getoption("--io-backend")
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
        backend = os.environ.get("HOLA_TEST_IO_BACKEND") or io_backend  # see `both_io_backends`
        if backend != "auto" and "--io-backend" not in list(args):
            args = list(args) + ["--io-backend", backend]
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


@pytest.fixture(params=["epoll", "uring"])
def both_io_backends(request, monkeypatch):
    """Runs the requesting test twice: once with `--io-backend epoll`, once with `uring` (which
    is epoll again on a kernel without a usable io_uring - the service says so on stderr).
    Modules that exercise the connection state machine opt in with
    ``pytestmark = pytest.mark.usefixtures("both_io_backends")``."""
    monkeypatch.setenv("HOLA_TEST_IO_BACKEND", request.param)
    return request.param


from hola_b200.fakes import FakeAloha, FakeZipkin  # noqa: E402,F401


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
