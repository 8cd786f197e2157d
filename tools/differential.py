#!/usr/bin/env python
"""Differential check of two running hola services.

SURVEY.md section 4 ("Differential vs reference"): the Java service cannot be run on the build
machines, so byte-equality with it is specification-derived, not observed. This tool is the
harness for the day a JVM and hola-swarm.jar are available: start both services with the same
environment and run

    python tools/differential.py http://127.0.0.1:8080 http://127.0.0.1:18080

It replays a fixed corpus of requests (every endpoint, methods, Accept negotiation, B3 headers,
unknown paths) against both base URLs and reports, per request, differences in status, body and
the headers whose values the service's own code determines. Headers that the container decides
(Date, Server, Connection, ...) are listed separately as "container-level" so that they can be
reviewed and, if wanted, matched in one place (append_response() in src/app/hola_app.cc).

Exit status: 0 when the contract-level comparison is clean, 1 otherwise.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path
from typing import Dict, List, Tuple
from urllib.parse import urlparse

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from hola_b200 import client  # noqa: E402

# (method, target, extra headers)
CORPUS: List[Tuple[str, str, Dict[str, str]]] = [
    ("GET", "/api/hola", {}),
    ("GET", "/api/hola?x=1", {}),
    ("GET", "/api/hola/", {}),
    ("HEAD", "/api/hola", {}),
    ("OPTIONS", "/api/hola", {}),
    ("POST", "/api/hola", {}),
    ("GET", "/api/hola", {"Accept": "text/plain"}),
    ("GET", "/api/hola", {"Accept": "application/json"}),
    ("GET", "/api/hola", {"Accept": "*/*;q=0.1"}),
    ("GET", "/api/hola", {"X-B3-TraceId": "463ac35c9f6413ad", "X-B3-SpanId": "a2fb4a1d1a96d312", "X-B3-Sampled": "1"}),
    ("GET", "/api/hola", {"X-B3-Sampled": "0"}),
    ("GET", "/api/health", {}),
    ("HEAD", "/api/health", {}),
    ("GET", "/api/hola-chaining", {}),
    ("GET", "/api/hola-chaining", {"Accept": "text/plain"}),
    ("GET", "/api/swagger.json", {}),
    ("GET", "/api/swagger.yaml", {}),
    ("GET", "/api/nothing-here", {}),
    ("GET", "/", {}),
    ("GET", "/index.html", {}),
    ("GET", "/webjars/swagger-ui/2.1.4/index.html", {}),
    ("GET", "/no/such/page", {}),
    ("DELETE", "/api/health", {}),
]

CONTRACT_HEADERS = ["content-type", "access-control-allow-origin", "allow", "content-length"]
CONTAINER_HEADERS = ["connection", "server", "x-powered-by", "transfer-encoding", "cache-control", "pragma"]


def fetch(base: str, method: str, target: str, headers: Dict[str, str]):
    u = urlparse(base)
    h = {"Host": "%s:%d" % (u.hostname, u.port or 80)}
    h.update(headers)
    with client.Connection(u.hostname, u.port or 80, timeout=10) as c:
        return c.request(method, target, headers=h)


def media_type(value):
    return None if value is None else value.split(";")[0].strip().lower()


def compare(a, b) -> Tuple[List[str], List[str]]:
    contract, container = [], []
    if a.status != b.status:
        contract.append("status %d != %d" % (a.status, b.status))
    if a.body != b.body:
        contract.append("body %r != %r" % (a.body[:120], b.body[:120]))
    for name in CONTRACT_HEADERS:
        va, vb = a.header(name), b.header(name)
        if name == "content-type":
            if media_type(va) != media_type(vb):
                contract.append("content-type media type %r != %r" % (va, vb))
            elif va != vb:
                container.append("content-type parameters %r != %r" % (va, vb))
        elif va != vb:
            contract.append("%s %r != %r" % (name, va, vb))
    for name in CONTAINER_HEADERS:
        va, vb = a.header(name), b.header(name)
        if va != vb:
            container.append("%s %r != %r" % (name, va, vb))
    names_a, names_b = [n.lower() for n in a.header_names()], [n.lower() for n in b.header_names()]
    if names_a != names_b:
        container.append("header order/set %r != %r" % (names_a, names_b))
    return contract, container


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("left", help="base URL of the first service, e.g. http://127.0.0.1:8080 (the reference)")
    ap.add_argument("right", help="base URL of the second service")
    ap.add_argument("--json", action="store_true", help="machine-readable report")
    args = ap.parse_args()
    report = []
    def own_authority(base: str) -> bytes:
        u = urlparse(base)
        return ("%s:%d" % (u.hostname, u.port or 80)).encode()

    for method, target, headers in CORPUS:
        a = fetch(args.left, method, target, headers)
        b = fetch(args.right, method, target, headers)
        # Error pages quote the request URL; each side's own host:port is not a difference.
        a.body = a.body.replace(own_authority(args.left), b"<authority>")
        b.body = b.body.replace(own_authority(args.right), b"<authority>")
        if a.header("Content-Length") != b.header("Content-Length") and a.body == b.body:
            a.headers = [(k, v) for k, v in a.headers if k.lower() != "content-length"]
            b.headers = [(k, v) for k, v in b.headers if k.lower() != "content-length"]
        contract, container = compare(a, b)
        report.append({"request": "%s %s %s" % (method, target, json.dumps(headers) if headers else ""),
                       "contract": contract, "container": container})
    bad = [r for r in report if r["contract"]]
    if args.json:
        print(json.dumps({"requests": len(report), "contract_differences": len(bad), "report": report}))
    else:
        for r in report:
            mark = "DIFF" if r["contract"] else ("note" if r["container"] else "same")
            print("[%s] %s" % (mark, r["request"].strip()))
            for line in r["contract"]:
                print("        contract:  " + line)
            for line in r["container"]:
                print("        container: " + line)
        print("%d request(s), %d with contract-level differences" % (len(report), len(bad)))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
