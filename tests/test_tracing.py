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


class ScriptedCollector:
    """A collector on a raw socket that answers every POST the way the test tells it to: which
    status, with or without a body, keep the connection or close it, or say nothing at all."""

    def __init__(self):
        import socket
        import threading
        self.sock = socket.socket()
        self.sock.bind(("127.0.0.1", 0))
        self.sock.listen(16)
        self.port = self.sock.getsockname()[1]
        self.url = "http://127.0.0.1:%d" % self.port
        self.reply = b"HTTP/1.1 202 Accepted\r\nContent-Length: 0\r\n\r\n"
        self.close_after_reply = False
        self.silent = False
        self.posts = []        # span lists, one per POST
        self.connections = 0
        self._stop = False
        self.thread = threading.Thread(target=self._serve, daemon=True)
        self.thread.start()

    def _serve(self):
        import threading
        self.sock.settimeout(0.1)
        while not self._stop:
            try:
                conn, _ = self.sock.accept()
            except OSError:
                continue
            self.connections += 1
            threading.Thread(target=self._handle, args=(conn,), daemon=True).start()

    def _handle(self, conn):
        conn.settimeout(5)
        buf = b""
        try:
            while not self._stop:
                while b"\r\n\r\n" not in buf:
                    chunk = conn.recv(65536)
                    if not chunk:
                        return
                    buf += chunk
                head, _, rest = buf.partition(b"\r\n\r\n")
                n = int([ln for ln in head.split(b"\r\n") if ln.lower().startswith(b"content-length")][0].split(b":")[1])
                while len(rest) < n:
                    rest += conn.recv(65536)
                self.posts.append(json.loads(rest[:n]))
                buf = rest[n:]
                if self.silent:
                    time.sleep(3)
                    return
                conn.sendall(self.reply)
                if self.close_after_reply:
                    return
        except OSError:
            pass
        finally:
            conn.close()

    def spans(self):
        return [s for post in self.posts for s in post]

    def stop(self):
        self._stop = True
        self.sock.close()


def _hit(srv, n):
    with client.Connection("127.0.0.1", srv.port) as c:
        for _ in range(n):
            assert c.request("GET", "/api/hola").status == 200


def _wait(cond, timeout=5.0):
    deadline = time.time() + timeout
    while time.time() < deadline and not cond():
        time.sleep(0.02)
    return cond()


def test_reporter_copes_with_every_kind_of_collector(hola_server):
    """The reporter keeps one connection to the collector and must survive whatever the collector
    does with it: answers with a body, closes after answering, closes while idle (the next POST
    meets a dead connection and is retried once on a fresh one), refuses the batch."""
    col = ScriptedCollector()
    try:
        srv = hola_server(env={"ZIPKIN_SERVER_URL": col.url}, args=["--zipkin-flush-ms", "40"])
        _hit(srv, 3)
        assert _wait(lambda: len(col.spans()) == 3) and col.connections == 1
        # A body on the answer is drained and the connection reused.
        col.reply = b"HTTP/1.1 200 OK\r\nContent-Length: 11\r\n\r\n{\"ok\":true}"
        _hit(srv, 2)
        assert _wait(lambda: len(col.spans()) == 5)
        _hit(srv, 1)
        assert _wait(lambda: len(col.spans()) == 6) and col.connections == 1
        # "Connection: close" on the answer: a new connection next time.
        col.reply = b"HTTP/1.1 202 Accepted\r\nConnection: close\r\nContent-Length: 0\r\n\r\n"
        col.close_after_reply = True
        _hit(srv, 1)
        assert _wait(lambda: len(col.spans()) == 7)
        _hit(srv, 1)
        assert _wait(lambda: len(col.spans()) == 8) and col.connections == 2
        # The collector drops the idle keep-alive connection without saying so: the next batch is
        # written into a dead socket, noticed, and sent again on a fresh connection - not lost.
        col.reply = b"HTTP/1.1 202 Accepted\r\nContent-Length: 0\r\n\r\n"
        _hit(srv, 1)
        assert _wait(lambda: len(col.spans()) == 9)
        time.sleep(0.2)
        _hit(srv, 2)
        assert _wait(lambda: len(col.spans()) >= 11)
        reported = srv.stats()["spans"]["reported"]
        assert reported == 11
        # A collector that refuses: the batch counts as dropped, requests are unaffected.
        col.close_after_reply = False
        col.reply = b"HTTP/1.1 500 Internal Server Error\r\nContent-Length: 4\r\n\r\nnope"
        _hit(srv, 4)
        assert _wait(lambda: srv.stats()["spans"]["dropped"] == 4)
        assert srv.stats()["spans"]["posts_failed"] >= 1 and srv.stats()["spans"]["reported"] == 11
        # An answer that is not HTTP at all.
        col.reply = b"SMTP ready\r\n\r\n"
        col.close_after_reply = True
        _hit(srv, 1)
        assert _wait(lambda: srv.stats()["spans"]["dropped"] == 5)
    finally:
        col.stop()


def test_small_ring_drops_at_the_producer_and_counts_it(hola_server):
    """--zipkin-ring: spans queued per worker before new ones are dropped (the reference's queue
    holds 1000 in total); --zipkin-max-batch: spans per POST."""
    col = ScriptedCollector()
    try:
        srv = hola_server(env={"ZIPKIN_SERVER_URL": col.url}, threads=1,
                          args=["--zipkin-flush-ms", "300", "--zipkin-ring", "8", "--zipkin-max-batch", "3"])
        with client.Connection("127.0.0.1", srv.port) as c:
            c.send_raw(client.build_request("GET", "/api/hola") * 50)  # one burst, far more than the ring holds
            for _ in range(50):
                assert c.read_response().status == 200
        assert _wait(lambda: len(col.spans()) == 8, timeout=3)
        assert all(len(post) <= 3 for post in col.posts) and len(col.posts) >= 3
        stats = srv.stats()["spans"]
        assert stats["reported"] == 8 and stats["dropped"] == 42
    finally:
        col.stop()


def test_shutdown_does_not_wait_for_a_collector_that_never_answers(hola_server):
    col = ScriptedCollector()
    col.silent = True
    try:
        srv = hola_server(env={"ZIPKIN_SERVER_URL": col.url}, args=["--zipkin-flush-ms", "60000"])
        _hit(srv, 2)
        t0 = time.time()
        assert srv.stop() == 0
        assert time.time() - t0 < 4.0  # the final flush gives up after about a second
        assert len(col.spans()) == 2   # it was sent; the answer never came
    finally:
        col.stop()


def test_collector_url_without_a_port_means_port_80(hola_server):
    srv = hola_server(env={"ZIPKIN_SERVER_URL": "http://127.0.0.1/zipkin/"}, args=["--zipkin-flush-ms", "50"])
    assert "tracing on" in srv.stderr
    _hit(srv, 2)
    assert _wait(lambda: srv.stats()["spans"]["dropped"] == 2)  # nothing listens on port 80 here
    srv6 = hola_server(env={"ZIPKIN_SERVER_URL": "http://[::1]"}, args=["--zipkin-flush-ms", "50"])
    assert "tracing on" in srv6.stderr
