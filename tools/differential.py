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

With ``--json`` the report also carries ``items``: one entry per decision that DESIGN.md section 4
records as taken from memory of a library that could not be run ([lib, unverified] in SURVEY.md) -
response-head shape, error bodies, JSON white space, Swagger text, metrics-stream fields, span
JSON, the X-B3-Sampled rule, java.util.Formatter corner cases, Feign's log lines - each with
``status`` same / different / not-compared and the evidence. Some of them can only be observed
when the tool starts the two services itself (their spans have to go to collectors the tool owns,
their stderr has to be captured, the greeting pattern has to be varied):

    python tools/differential.py --launch-left "java -Dswarm.http.port={port} -jar hola-swarm.jar" \\
                                 --launch-right "build/hola --port {port} --reference-logging" --json

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


def read_stream_events(base: str, count: int = 2, timeout: float = 5.0):
    """First `count` data events of /hystrix.stream plus the response's content type."""
    u = urlparse(base)
    with client.Connection(u.hostname, u.port or 80, timeout=timeout) as c:
        c.send_raw(client.build_request("GET", "/hystrix.stream", {"Host": "%s:%d" % (u.hostname, u.port or 80)}))
        while b"\r\n\r\n" not in c.buf:
            if not c._fill():
                return None, []
        head, _, c.buf = c.buf.partition(b"\r\n\r\n")
        ctype = None
        for line in head.split(b"\r\n")[1:]:
            k, _, v = line.partition(b":")
            if k.strip().lower() == b"content-type":
                ctype = v.strip().decode("latin-1")
        chunked = b"chunked" in head.lower()
        data, events = b"", []
        try:
            while len(events) < count:
                if chunked:
                    piece = c.read_chunk()
                    if piece is None:
                        break
                    data += piece
                else:
                    if not c._fill():
                        break
                    data, c.buf = data + c.buf, b""
                while b"\n\n" in data:
                    ev, _, data = data.partition(b"\n\n")
                    if ev.startswith(b"data:"):
                        events.append(json.loads(ev[5:].strip()))
        except (OSError, ValueError):
            pass
        return ctype, events


def item(ident: str, what: str, where: str, same, details) -> Dict:
    status = "not-compared" if same is None else ("same" if same else "different")
    return {"id": ident, "what": what, "decided_in": where, "status": status, "details": details}


def unverified_items(left: str, right: str, report: List[Dict], raw: Dict) -> List[Dict]:
    """The URL-only part of the [lib, unverified] checklist."""
    items = []
    container = [(r["request"].strip(), r["container"]) for r in report if r["container"]]
    items.append(item("response-head", "header set and order, Connection, Server/X-Powered-By, Content-Type parameters",
                      "append_response() in src/app/hola_app.cc", not container, container))
    errors = [(r["request"].strip(), r["contract"]) for r in report
              if raw[r["request"]][0].status >= 400 and any(c.startswith("body") for c in r["contract"])]
    statuses = sorted({raw[r["request"]][0].status for r in report if raw[r["request"]][0].status >= 400})
    items.append(item("error-bodies", "bodies of the %s answers (RESTEasy messages, Undertow error page)" % "/".join(map(str, statuses)),
                      "HolaApp::handle() in src/app/hola_app.cc", not errors, errors))
    for ident, what, where, match in (
            ("chaining-json-whitespace", "byte-exact body of GET /api/hola-chaining", "json_encode_string_list() in src/core/json.cc", "/api/hola-chaining "),
            ("swagger-text", "byte-exact /api/swagger.json and /api/swagger.yaml", "src/app/static_content.cc", "/api/swagger")):
        hits = [(r["request"].strip(), r["contract"]) for r in report if match in r["request"] + " " and r["contract"]]
        items.append(item(ident, what, where, not hits, hits))
    try:
        # A command has to have run for the stream to carry data events.
        for base in (left, right):
            fetch(base, "GET", "/api/hola-chaining", {})
        import time
        time.sleep(1.2)
        (ct_a, ev_a), (ct_b, ev_b) = read_stream_events(left), read_stream_events(right)
        details = []
        if ct_a != ct_b:
            details.append("content-type %r != %r" % (ct_a, ct_b))
        if len(ev_a) != len(ev_b):
            details.append("%d data events != %d within the time allowed" % (len(ev_a), len(ev_b)))
        for a, b in zip(ev_a, ev_b):
            if set(a) != set(b):
                details.append({"type": a.get("type"), "only_left": sorted(set(a) - set(b)), "only_right": sorted(set(b) - set(a))})
            static = [k for k in a if k in b and k.startswith("propertyValue_") and a[k] != b[k]]
            if static:
                details.append({"type": a.get("type"), "different_property_values": {k: [a[k], b[k]] for k in static}})
        items.append(item("hystrix-stream-fields", "field names and property values of the /hystrix.stream payloads",
                          "hystrix_command_json()/hystrix_pool_json() in src/core/breaker.cc", not details, details))
    except Exception as exc:  # noqa: BLE001 - one side may not have the endpoint at all
        items.append(item("hystrix-stream-fields", "field names of the /hystrix.stream payloads", "src/core/breaker.cc", None,
                          ["could not read the stream: %s" % exc]))
    return items


def launched_items(launch_left: str, launch_right: str) -> List[Dict]:
    """The part of the checklist that needs both services started by this tool."""
    import shlex
    import time
    from hola_b200.fakes import FakeZipkin
    from hola_b200.reference import ReferenceServer
    items = []

    def start(command: str, extra_env=None, pattern=None):
        zipkin = FakeZipkin()
        argv = shlex.split(command)
        if pattern is not None:
            argv = argv[:1] + ["-Dhello=" + pattern] + argv[1:]
        env = {"ZIPKIN_SERVER_URL": zipkin.url, "HOSTNAME": None}
        env.update(extra_env or {})
        srv = ReferenceServer(argv, env=env, ready_path="/api/health")
        srv.keep_log = True
        srv.start()
        return srv, zipkin

    def normal_span(s):
        out = {"has_parent": "parentId" in s, "name": s.get("name"), "id_width": len(s.get("id", "")), "trace_width": len(s.get("traceId", "")),
               "annotations": [a.get("value") for a in s.get("annotations", [])],
               "endpoint_keys": sorted((s.get("annotations") or [{}])[0].get("endpoint", {}).keys()),
               "tags": sorted(t.get("key") for t in s.get("binaryAnnotations", [])),
               "keys": sorted(s.keys())}
        return out

    a, za = start(launch_left)
    b, zb = start(launch_right)
    try:
        # Span JSON and the X-B3-Sampled rule.
        probes = [({}, "new trace"), ({"X-B3-TraceId": "463ac35c9f6413ad", "X-B3-SpanId": "a2fb4a1d1a96d312", "X-B3-Sampled": "1"}, "joined"),
                  ({"X-B3-TraceId": "463ac35c9f6413ad", "X-B3-SpanId": "a2fb4a1d1a96d312"}, "ids without X-B3-Sampled"),
                  ({"X-B3-Sampled": "0"}, "not sampled")]
        for headers, _ in probes:
            fetch(a.base_url, "GET", "/api/hola", headers)
            fetch(b.base_url, "GET", "/api/hola", headers)
        sa, sb = za.wait_for_spans(3, timeout=4.0), zb.wait_for_spans(3, timeout=4.0)
        details = []
        if len(sa) != len(sb):
            details.append("%d spans reported != %d (four requests, one of them unsampled)" % (len(sa), len(sb)))
        # Spans of different worker threads reach the collector in no particular order.
        na = sorted((normal_span(x) for x in sa), key=lambda n: json.dumps(n, sort_keys=True))
        nb = sorted((normal_span(y) for y in sb), key=lambda n: json.dumps(n, sort_keys=True))
        for nx, ny in zip(na, nb):
            if nx != ny:
                details.append({"left": nx, "right": ny})
        items.append(item("zipkin-span-json", "shape of the Zipkin v1 span JSON (name, id width, annotations, tags)", "src/core/trace.cc",
                          not details, details))
        joined = lambda spans: sorted(s["traceId"] for s in spans if s.get("traceId", "").endswith("463ac35c9f6413ad"))  # noqa: E731
        items.append(item("b3-ids-need-sampled-header", "caller's ids are honoured only when X-B3-Sampled is present (Brave 3.9 adapter)",
                          "start_server_trace() in src/core/trace.cc", len(joined(sa)) == len(joined(sb)),
                          {"left_spans_with_callers_trace_id": len(joined(sa)), "right": len(joined(sb))}))
        # Feign's log lines.
        for base in (a.base_url, b.base_url):
            fetch(base, "GET", "/api/hola-chaining", {})
        time.sleep(0.3)
        import re
        feign = lambda log: [re.sub(r"\(\d+ms\)", "(Nms)", ln.strip()) for ln in log.splitlines() if "[AlohaService#aloha]" in ln]  # noqa: E731
        la, lb = feign(a.log_tail(20000)), feign(b.log_tail(20000))
        items.append(item("feign-basic-log", "Feign Logger.Level.BASIC lines on stderr per downstream call", "HolaApp::send_chain_attempt()",
                          la == lb, {"left": la[:6], "right": lb[:6]}))
    finally:
        for x in (a, b, za, zb):
            x.stop()
    # java.util.Formatter corner cases: one launch pair per pattern (-Dhello=... right after the program name).
    details, compared = [], True
    for pattern in ("Hola %% de %s", "[%-12s]", "[%12.3s]", "%5%|%s", "%S", "%s %s", "%d"):
        started = []
        try:
            for command in (launch_left, launch_right):
                started.extend(start(command, pattern=pattern))
            a, b = started[0], started[2]
            ra, rb = fetch(a.base_url, "GET", "/api/hola", {}), fetch(b.base_url, "GET", "/api/hola", {})
            if (ra.status, ra.body if ra.status == 200 else b"") != (rb.status, rb.body if rb.status == 200 else b""):
                details.append({"pattern": pattern, "left": [ra.status, ra.body[:80].decode("latin-1")], "right": [rb.status, rb.body[:80].decode("latin-1")]})
        except Exception as exc:  # noqa: BLE001 - a side that does not take -Dhello=... cannot be compared
            details.append({"pattern": pattern, "error": str(exc)[:200]})
            compared = False
        finally:
            for x in started:
                x.stop()
        if not compared:
            break
    items.append(item("formatter-corner-cases", "String.format(pattern, hostname) for patterns beyond a single %s (JDK 8: %% ignores width)",
                      "src/core/formatter.cc", (not details) if compared else None, details))
    return items


NOT_COVERED = [
    item("dns-cache-ttl", "JVM-style DNS cache (30 s positive / 10 s negative)", "src/net/resolver.cc", None,
         ["timing behaviour against a name server: needs a controllable DNS in front of both services"]),
    item("feign-retryer-schedule", "Feign default Retryer pauses (150/225/337/506 ms) inside one Hystrix command", "HolaApp::send_chain_attempt()", None,
         ["timing behaviour against a failing downstream: compare with tests/test_chaining.py::test_connection_failures_are_retried_like_feign"]),
]


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("left", nargs="?", help="base URL of the first service, e.g. http://127.0.0.1:8080 (the reference)")
    ap.add_argument("right", nargs="?", help="base URL of the second service")
    ap.add_argument("--launch-left", help="command that starts the first service; {port} is replaced by a free port")
    ap.add_argument("--launch-right", help="command that starts the second service")
    ap.add_argument("--json", action="store_true", help="machine-readable report")
    args = ap.parse_args()
    launched = []
    if args.launch_left and args.launch_right:
        import shlex
        from hola_b200.reference import ReferenceServer
        env = {"ZIPKIN_SERVER_URL": "http://127.0.0.1:9", "HOSTNAME": None}
        launched = [ReferenceServer(shlex.split(args.launch_left), env=env, ready_path="/api/health").start(),
                    ReferenceServer(shlex.split(args.launch_right), env=env, ready_path="/api/health").start()]
        args.left, args.right = launched[0].base_url, launched[1].base_url
    if not args.left or not args.right:
        ap.error("give two base URLs, or --launch-left and --launch-right")
    report = []
    raw = {}
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
        key = "%s %s %s" % (method, target, json.dumps(headers) if headers else "")
        raw[key] = (a, b)
        report.append({"request": key, "contract": contract, "container": container})
    bad = [r for r in report if r["contract"]]
    items = unverified_items(args.left, args.right, report, raw) if args.json else []
    for srv in launched:
        srv.stop()
    if args.json and args.launch_left and args.launch_right:
        try:
            items += launched_items(args.launch_left, args.launch_right)
        except Exception as exc:  # noqa: BLE001 - report what could be compared
            items.append(item("launched-comparisons", "span JSON, B3 rule, Feign log, formatter", "-", None, ["aborted: %s" % str(exc)[:300]]))
    elif args.json:
        items += [item(i, w, d, None, ["needs --launch-left/--launch-right: the tool must own the collectors, the logs and the greeting pattern"])
                  for i, w, d in (("zipkin-span-json", "shape of the Zipkin v1 span JSON", "src/core/trace.cc"),
                                  ("b3-ids-need-sampled-header", "caller's ids honoured only with X-B3-Sampled", "src/core/trace.cc"),
                                  ("feign-basic-log", "Feign BASIC log lines", "HolaApp::send_chain_attempt()"),
                                  ("formatter-corner-cases", "String.format corner cases", "src/core/formatter.cc"))]
    if args.json:
        items += NOT_COVERED
        print(json.dumps({"requests": len(report), "contract_differences": len(bad), "report": report, "items": items}))
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
