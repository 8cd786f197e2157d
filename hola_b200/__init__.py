"""hola_b200 - Python side of the hola rebuild: build, launch, load-test and poke the native service.

The service itself is native code (C++17, ``src/``); the reference it rebuilds is a Java EE
microservice (``/root/reference/src/main/java/com/redhat/developers/msa/hola/``). This package
is the tooling around it:

* :func:`build` / :func:`paths`      compile the executables and ``libhola.so`` in-tree (``build/``)
* :class:`hola_b200.server.HolaServer`   run ``build/hola`` as a child process
* :class:`hola_b200.native.Service`      run the service in-process through the C ABI
* :mod:`hola_b200.native`                ctypes bindings to the building blocks (parsers, breaker, ...)
* :func:`hola_b200.loadgen.run`          drive ``build/hola-load`` and parse its JSON report
* :mod:`hola_b200.client`                a tiny raw-socket HTTP client for byte-exact assertions
* :mod:`hola_b200.fakes`                 fake aloha / Zipkin / DNS servers for tests and manual trials
* :mod:`hola_b200.reference`             launcher for the Java reference (or any stand-in) used by ``bench.py --impl reference``

There are deliberately no ``models/``, ``ops/`` or ``parallel/`` sub-packages: the reference has
no model, tensor op, kernel or collective, and none was invented.
"""
from __future__ import annotations

import os
import subprocess
from dataclasses import dataclass
from pathlib import Path

__version__ = "0.0.1"

REPO_ROOT = Path(__file__).resolve().parent.parent
BUILD_DIR = REPO_ROOT / "build"


@dataclass(frozen=True)
class Paths:
    hola: Path
    hola_load: Path
    libhola: Path
    unit_tests: Path

    def all_exist(self) -> bool:
        return all(p.exists() for p in (self.hola, self.hola_load, self.libhola, self.unit_tests))


def paths(variant: str = "release") -> Paths:
    """Build outputs for a variant: release, asan, tsan, or static (the fully static executable
    that goes into the FROM-scratch image; the other artefacts are the release ones)."""
    if variant == "static":
        return Paths(BUILD_DIR / "hola-static", BUILD_DIR / "hola-load", BUILD_DIR / "libhola.so", BUILD_DIR / "hola_tests")
    base = BUILD_DIR if variant == "release" else BUILD_DIR / variant
    return Paths(base / "hola", base / "hola-load", base / "libhola.so", base / "hola_tests")


def _sources_newer_than(target: Path) -> bool:
    if not target.exists():
        return True
    stamp = target.stat().st_mtime
    for sub in ("src", "include", "tests/cpp"):
        for path in (REPO_ROOT / sub).rglob("*"):
            if path.is_file() and path.stat().st_mtime > stamp:
                return True
    return (REPO_ROOT / "Makefile").stat().st_mtime > stamp


def build(variant: str = "release", force: bool = False, quiet: bool = True) -> Paths:
    """Compile everything with ``make`` (a no-op when the outputs are newer than the sources)."""
    p = paths(variant)
    if not force and p.all_exist() and not any(_sources_newer_than(t) for t in (p.hola, p.hola_load, p.libhola, p.unit_tests)):
        return p
    jobs = str(max(1, len(os.sched_getaffinity(0))))
    if variant == "static":
        cmd = ["make", "-C", str(REPO_ROOT), "-j", jobs, "all", "static"]
    elif variant == "release":
        cmd = ["make", "-C", str(REPO_ROOT), "-j", jobs]
    else:
        cmd = ["make", "-C", str(REPO_ROOT), "-j", jobs, variant]  # asan / tsan pick their own compiler
    result = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if result.returncode != 0:
        raise RuntimeError("native build failed:\n" + result.stdout[-4000:])
    if not quiet:
        print(result.stdout)
    if not p.all_exist():
        raise RuntimeError("native build finished but outputs are missing: %r" % (p,))
    return p
