"""GET /hystrix.stream (HystrixServlet.java:12-17): server-sent events with Hystrix 1.5.1's
field names, so that a Hystrix dashboard / Turbine can consume it (pom.xml:82)."""
from __future__ import annotations

import json
import time

from hola_b200 import client


def open_stream(srv, query=""):
    c = client.Connection("127.0.0.1", srv.port, timeout=5)
    c.send_raw(client.build_request("GET", "/hystrix.stream" + query))
    # Head only; the body never ends.
    while b"\r\n\r\n" not in c.buf:
        assert c._fill()
    head, _, c.buf = c.buf.partition(b"\r\n\r\n")
    return c, head.decode()


def read_events(c, n):
    events = []
    pending = b""
    while len(events) < n:
        chunk = c.read_chunk()
        assert chunk is not None
        pending += chunk
        while b"\n\n" in pending:
            ev, _, pending = pending.partition(b"\n\n")
            events.append(ev.decode())
    return events


def test_stream_pings_until_a_command_has_run(hola_server):
    srv = hola_server()
    c, head = open_stream(srv, "?delay=50")
    assert head.startswith("HTTP/1.1 200 OK")
    assert "Content-Type: text/event-stream;charset=UTF-8" in head
    assert "Transfer-Encoding: chunked" in head
    assert "no-cache" in head
    assert "Access-Control-Allow-Origin" not in head  # a servlet, not a JAX-RS resource: no CORS filter
    events = read_events(c, 3)
    assert events == ["ping: "] * 3
    c.close()


def test_stream_for_an_http10_client_is_not_chunked(hola_server):
    srv = hola_server()
    c = client.Connection("127.0.0.1", srv.port, timeout=5)
    c.send_raw(client.build_request("GET", "/hystrix.stream?delay=30", version="HTTP/1.0"))
    data = b""
    while data.count(b"ping: \n\n") < 3:
        assert c._fill()
        data = c.buf
    head, _, body = data.partition(b"\r\n\r\n")
    assert b"Transfer-Encoding" not in head and b"Connection: close" in head
    assert body.startswith(b"ping: \n\nping: \n\n")  # raw events, no chunk-size lines
    c.close()


def test_stream_reports_command_and_pool_after_calls(hola_server, fake_aloha):
    srv = hola_server(args=["--aloha-url", fake_aloha.url], threads=1)
    for _ in range(3):
        client.get("127.0.0.1", srv.port, "/api/hola-chaining")
    fake_aloha.mode = "500"
    client.get("127.0.0.1", srv.port, "/api/hola-chaining")
    time.sleep(1.1)  # rolling counters publish closed one-second buckets
    c, _ = open_stream(srv, "?delay=100")
    events = read_events(c, 2)
    c.close()
    assert all(e.startswith("data: ") for e in events)
    cmd = json.loads(events[0][6:])
    pool = json.loads(events[1][6:])
    assert cmd["type"] == "HystrixCommand" and pool["type"] == "HystrixThreadPool"
    assert cmd["name"] == "AlohaService#aloha()"
    assert cmd["group"] == fake_aloha.url and cmd["threadPool"] == pool["name"]
    assert cmd["rollingCountSuccess"] == 3 and cmd["rollingCountFailure"] == 1
    assert cmd["rollingCountFallbackSuccess"] == 1
    assert cmd["requestCount"] == 4 and cmd["errorCount"] == 1 and cmd["errorPercentage"] == 25
    assert cmd["isCircuitBreakerOpen"] is False
    assert cmd["propertyValue_executionIsolationThreadTimeoutInMilliseconds"] == 600
    assert cmd["propertyValue_circuitBreakerRequestVolumeThreshold"] == 5
    assert set(cmd["latencyExecute"]) == {"0", "25", "50", "75", "90", "95", "99", "99.5", "100"}
    assert abs(cmd["currentTime"] / 1000 - time.time()) < 10
    assert pool["currentCorePoolSize"] == 10 and pool["rollingCountThreadsExecuted"] == 4
    expected_cmd_fields = {
        "type", "name", "group", "currentTime", "isCircuitBreakerOpen", "errorPercentage", "errorCount", "requestCount",
        "rollingCountBadRequests", "rollingCountCollapsedRequests", "rollingCountEmit", "rollingCountExceptionsThrown",
        "rollingCountFailure", "rollingCountFallbackEmit", "rollingCountFallbackFailure", "rollingCountFallbackMissing",
        "rollingCountFallbackRejection", "rollingCountFallbackSuccess", "rollingCountResponsesFromCache",
        "rollingCountSemaphoreRejected", "rollingCountShortCircuited", "rollingCountSuccess", "rollingCountThreadPoolRejected",
        "rollingCountTimeout", "currentConcurrentExecutionCount", "rollingMaxConcurrentExecutionCount", "latencyExecute_mean",
        "latencyExecute", "latencyTotal_mean", "latencyTotal", "propertyValue_circuitBreakerRequestVolumeThreshold",
        "propertyValue_circuitBreakerSleepWindowInMilliseconds", "propertyValue_circuitBreakerErrorThresholdPercentage",
        "propertyValue_circuitBreakerForceOpen", "propertyValue_circuitBreakerForceClosed", "propertyValue_circuitBreakerEnabled",
        "propertyValue_executionIsolationStrategy", "propertyValue_executionIsolationThreadTimeoutInMilliseconds",
        "propertyValue_executionTimeoutInMilliseconds", "propertyValue_executionIsolationThreadInterruptOnTimeout",
        "propertyValue_executionIsolationThreadPoolKeyOverride", "propertyValue_executionIsolationSemaphoreMaxConcurrentRequests",
        "propertyValue_fallbackIsolationSemaphoreMaxConcurrentRequests", "propertyValue_metricsRollingStatisticalWindowInMilliseconds",
        "propertyValue_requestCacheEnabled", "propertyValue_requestLogEnabled", "reportingHosts", "threadPool"}
    assert set(cmd) == expected_cmd_fields
    assert set(pool) == {
        "type", "name", "currentTime", "currentActiveCount", "currentCompletedTaskCount", "currentCorePoolSize",
        "currentLargestPoolSize", "currentMaximumPoolSize", "currentPoolSize", "currentQueueSize", "currentTaskCount",
        "rollingCountThreadsExecuted", "rollingMaxActiveThreads", "rollingCountCommandRejections",
        "propertyValue_queueSizeRejectionThreshold", "propertyValue_metricsRollingStatisticalWindowInMilliseconds", "reportingHosts"}


def test_stream_shows_open_circuit(hola_server):
    srv = hola_server(args=["--aloha-url", "http://127.0.0.1:9/", "--feign-max-attempts", "1"], threads=1)
    for _ in range(6):
        client.get("127.0.0.1", srv.port, "/api/hola-chaining")
    time.sleep(0.6)
    client.get("127.0.0.1", srv.port, "/api/hola-chaining")  # short-circuited
    time.sleep(1.1)
    c, _ = open_stream(srv, "?delay=100")
    cmd = json.loads(read_events(c, 1)[0][6:])
    c.close()
    assert cmd["isCircuitBreakerOpen"] is True
    assert cmd["rollingCountShortCircuited"] >= 1 and cmd["errorPercentage"] == 100


def test_stream_connection_limit_and_release(hola_server):
    """hystrix.stream.maxConcurrentConnections defaults to 5; the 6th client gets a 503."""
    srv = hola_server(threads=2)
    streams = [open_stream(srv, "?delay=50")[0] for _ in range(5)]
    r = client.get("127.0.0.1", srv.port, "/hystrix.stream")
    assert r.status == 503 and b"MaxConcurrentConnections reached: 5" in r.body
    streams[0].close()
    deadline = time.time() + 3
    ok = False
    while time.time() < deadline and not ok:
        time.sleep(0.1)
        c = client.Connection("127.0.0.1", srv.port)
        c.send_raw(client.build_request("GET", "/hystrix.stream?delay=50"))
        while b"\r\n\r\n" not in c.buf:
            c._fill()
        ok = c.buf.startswith(b"HTTP/1.1 200")
        c.close()
    assert ok
    for s in streams[1:]:
        s.close()
    # Other traffic is unaffected by open streams, and shutdown does not hang on them.
    s = open_stream(srv)[0]
    assert client.get("127.0.0.1", srv.port, "/api/hola").status == 200
    assert srv.stop() == 0
    s.close()


def test_counters_of_all_workers_are_merged(hola_server, fake_aloha):
    """The breaker's counters are kept per event-loop thread and merged when somebody looks
    (core/breaker.h). Requests spread over eight workers must add up exactly, and failures seen by
    different workers must trip the one circuit."""
    srv = hola_server(args=["--aloha-url", fake_aloha.url, "--feign-max-attempts", "1"], threads=8)
    conns = [client.Connection("127.0.0.1", srv.port) for _ in range(16)]  # two per worker (balanced placement)
    try:
        for c in conns:
            for _ in range(2):
                assert c.request("GET", "/api/hola-chaining").status == 200
        stats = srv.stats()
        assert [w["requests"] for w in stats["workers"]] == [4] * 8
        time.sleep(1.1)
        s, _ = open_stream(srv, "?delay=100")
        events = read_events(s, 2)
        cmd, pool = json.loads(events[0][6:]), json.loads(events[1][6:])
        assert cmd["rollingCountSuccess"] == 32 and cmd["requestCount"] == 32 and cmd["errorCount"] == 0
        assert pool["rollingCountThreadsExecuted"] == 32 and pool["currentCompletedTaskCount"] == 32
        # One failure on each of six workers: no worker alone has seen five, together they have.
        fake_aloha.mode = "500"
        for c in conns[:12:2]:
            assert json.loads(c.request("GET", "/api/hola-chaining").body)[1] == "Aloha response (fallback)"
        fake_aloha.mode = "ok"
        for c in conns[:6]:
            c.request("GET", "/api/hola-chaining")
        time.sleep(0.6)
        s.close()  # events have piled up in the socket meanwhile: look through a fresh subscription
        s, _ = open_stream(srv, "?delay=100")
        cmd = json.loads(read_events(s, 1)[0][6:])
        assert cmd["errorCount"] == 6 and cmd["isCircuitBreakerOpen"] is False  # 6 of 44: below 50 %
        fake_aloha.mode = "500"
        for _ in range(5):
            for c in conns[:16:2]:
                c.request("GET", "/api/hola-chaining")
        time.sleep(0.6)
        # The next call evaluates the closed buckets: 46 errors of 84 requests -> open, for every worker.
        for c in conns:
            assert json.loads(c.request("GET", "/api/hola-chaining").body) == ["Hola de unknown", "Aloha response (fallback)"]
        n_before = len(fake_aloha.snapshot())
        for c in conns:
            c.request("GET", "/api/hola-chaining")
        assert len(fake_aloha.snapshot()) == n_before  # short-circuited on all eight workers: nothing reaches aloha
        time.sleep(1.1)
        s.close()
        s, _ = open_stream(srv, "?delay=100")
        cmd = json.loads(read_events(s, 1)[0][6:])
        assert cmd["isCircuitBreakerOpen"] is True and cmd["rollingCountShortCircuited"] >= 16
        s.close()
    finally:
        for c in conns:
            c.close()
