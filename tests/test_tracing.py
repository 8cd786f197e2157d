"""Zipkin tracing (ZipKinFilter.java:38-44, ZipkinResource.java:33-41, HolaResource.java:97-108),
tested against a fake collector and a fake downstream."""
from __future__ import annotations

import json
import time

from hola_b200 import client


def by_kind(spans):
    server = [s for s in spans if s["annotations"][0]["value"] == "sr"]
    clients = [s for s in spans if s["annotations"][0]["value"] == "cs"]
    return server, clients


def tag(span, key):
    for t in span["binaryAnnotations"]:
        if t["key"] == key:
            return t["value"]
    return None


def test_server_span_for_hola(hola_server, fake_zipkin):
    srv = hola_server(env={"ZIPKIN_SERVER_URL": fake_zipkin.url}, args=["--zipkin-flush-ms", "100"])
    assert "tracing on" in srv.stderr
    r = client.get("127.0.0.1", srv.port, "/api/hola?who=me", headers={"Host": "localhost:8080"})
    assert r.status == 200
    spans = fake_zipkin.wait_for_spans(1)
    assert len(spans) == 1
    s = spans[0]
    post = fake_zipkin.snapshot()[0]
    assert post["path"] == "/api/v1/spans" and post["content_type"] == "application/json"  # Zipkin v1 endpoint
    assert s["traceId"] == s["id"] and len(s["id"]) == 16 and "parentId" not in s  # new root trace
    assert s["name"] == "get"
    assert [a["value"] for a in s["annotations"]] == ["sr", "ss"]
    assert all(a["endpoint"]["serviceName"] == "hola" for a in s["annotations"])  # ZipkinResource.java:36
    assert tag(s, "http.url") == "http://localhost:8080/api/hola?who=me"
    assert tag(s, "http.status_code") is None  # only reported for non-2xx
    assert s["duration"] >= 1 and abs(s["timestamp"] / 1e6 - time.time()) < 10


def test_only_the_two_mapped_urls_are_traced(hola_server, fake_zipkin):
    """Health is deliberately left out (ZipKinFilter.java:42-43); so is everything else."""
    srv = hola_server(env={"ZIPKIN_SERVER_URL": fake_zipkin.url}, args=["--zipkin-flush-ms", "100", "--aloha-url", "http://127.0.0.1:9/"])
    for target in ("/api/health", "/", "/api/swagger.json", "/api/hola/", "/api/nothing"):
        client.get("127.0.0.1", srv.port, target)
    client.get("127.0.0.1", srv.port, "/api/hola")
    client.get("127.0.0.1", srv.port, "/api/hola-chaining")
    spans = fake_zipkin.wait_for_spans(2)
    time.sleep(0.3)
    spans = fake_zipkin.spans()
    urls = sorted(tag(s, "http.url").split("localhost")[1] for s in spans)
    assert urls == ["/api/hola", "/api/hola-chaining"]


def test_inbound_b3_headers_are_joined(hola_server, fake_zipkin):
    srv = hola_server(env={"ZIPKIN_SERVER_URL": fake_zipkin.url}, args=["--zipkin-flush-ms", "100"])
    client.get("127.0.0.1", srv.port, "/api/hola", headers={
        "X-B3-TraceId": "463ac35c9f6413ad", "X-B3-SpanId": "a2fb4a1d1a96d312", "X-B3-ParentSpanId": "0020000000000001",
        "X-B3-Sampled": "1"})
    s = fake_zipkin.wait_for_spans(1)[0]
    assert (s["traceId"], s["id"], s["parentId"]) == ("463ac35c9f6413ad", "a2fb4a1d1a96d312", "0020000000000001")
    assert "timestamp" not in s  # a span shared with the caller carries annotations only


def test_not_sampled_requests_are_not_reported(hola_server, fake_zipkin, fake_aloha):
    srv = hola_server(env={"ZIPKIN_SERVER_URL": fake_zipkin.url}, args=["--zipkin-flush-ms", "100", "--aloha-url", fake_aloha.url])
    client.get("127.0.0.1", srv.port, "/api/hola-chaining", headers={"X-B3-Sampled": "0"})
    time.sleep(0.4)
    assert fake_zipkin.spans() == []
    down = fake_aloha.snapshot()[0]["headers"]
    assert down.get("x-b3-sampled") == "0" and "x-b3-traceid" not in down


def test_chaining_propagates_a_child_span_downstream(hola_server, fake_zipkin, fake_aloha):
    srv = hola_server(env={"ZIPKIN_SERVER_URL": fake_zipkin.url}, args=["--zipkin-flush-ms", "100", "--aloha-url", fake_aloha.url])
    r = client.get("127.0.0.1", srv.port, "/api/hola-chaining")
    assert json.loads(r.body)[1:] == fake_aloha.payload
    spans = fake_zipkin.wait_for_spans(2)
    server, clients = by_kind(spans)
    assert len(server) == 1 and len(clients) == 1
    s, c = server[0], clients[0]
    assert c["traceId"] == s["traceId"] and c["parentId"] == s["id"] and c["id"] != s["id"]
    assert [a["value"] for a in c["annotations"]] == ["cs", "cr"]
    assert tag(c, "http.url") == fake_aloha.url + "api/aloha-chaining"
    down = fake_aloha.snapshot()[0]["headers"]  # B3 headers on the wire, HolaResource.java:100
    assert down["x-b3-traceid"] == s["traceId"]
    assert down["x-b3-spanid"] == c["id"]
    assert down["x-b3-parentspanid"] == s["id"]
    assert down["x-b3-sampled"] == "1"
    # The client span sits inside the server span.
    assert s["timestamp"] <= c["timestamp"] and c["timestamp"] + c["duration"] <= s["timestamp"] + s["duration"] + 1000


def test_downstream_error_status_is_tagged(hola_server, fake_zipkin, fake_aloha):
    fake_aloha.mode = "500"
    srv = hola_server(env={"ZIPKIN_SERVER_URL": fake_zipkin.url}, args=["--zipkin-flush-ms", "100", "--aloha-url", fake_aloha.url])
    client.get("127.0.0.1", srv.port, "/api/hola-chaining")
    _, clients = by_kind(fake_zipkin.wait_for_spans(2))
    assert tag(clients[0], "http.status_code") == "500"


def test_dead_collector_never_hurts_requests(hola_server):
    """Tracing must be harmless when Zipkin is down: spans are dropped, requests fly."""
    srv = hola_server(env={"ZIPKIN_SERVER_URL": "http://127.0.0.1:9"}, args=["--zipkin-flush-ms", "50"])
    with client.Connection("127.0.0.1", srv.port) as c:
        t0 = time.time()
        for _ in range(3000):  # more than the ring holds
            assert c.request("GET", "/api/hola").status == 200
        assert time.time() - t0 < 10
    assert srv.stop() == 0


def test_unset_url_disables_tracing_instead_of_crashing(hola_server, fake_aloha):
    """Deliberate improvement over the reference, which cannot deploy without the variable."""
    srv = hola_server(args=["--aloha-url", fake_aloha.url])
    assert "tracing on" not in srv.stderr
    assert client.get("127.0.0.1", srv.port, "/api/hola-chaining").status == 200
    down = fake_aloha.snapshot()[0]["headers"]
    assert not any(k.startswith("x-b3-") for k in down)


def test_bad_url_is_reported_and_ignored(hola_server):
    srv = hola_server(env={"ZIPKIN_SERVER_URL": "ftp://nope"})
    assert "tracing disabled" in srv.stderr
    assert client.get("127.0.0.1", srv.port, "/api/hola").status == 200


def test_spans_flushed_on_shutdown(hola_server, fake_zipkin):
    srv = hola_server(env={"ZIPKIN_SERVER_URL": fake_zipkin.url}, args=["--zipkin-flush-ms", "60000"])
    for _ in range(7):
        client.get("127.0.0.1", srv.port, "/api/hola")
    assert fake_zipkin.spans() == []
    srv.stop()
    assert len(fake_zipkin.spans()) == 7
