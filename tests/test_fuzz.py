"""Socket-level robustness: whatever bytes arrive, the server neither crashes nor wedges, and it
never answers a request it should not have understood. Run it with ``--variant asan`` to have
AddressSanitizer/UBSan watch the same traffic."""
from __future__ import annotations

import random
import socket

from hola_b200 import client

GOOD = client.build_request("GET", "/api/hola")
SEEDS = [
    GOOD,
    client.build_request("GET", "/api/hola-chaining"),
    client.build_request("POST", "/api/health", body=b"x" * 37),
    b"POST /api/hola HTTP/1.1\r\nHost: x\r\nTransfer-Encoding: chunked\r\n\r\n5\r\nhello\r\n0\r\n\r\n",
    client.build_request("GET", "/hystrix.stream?delay=10"),
    client.build_request("GET", "/api/hola", headers={"X-B3-TraceId": "f" * 16, "X-B3-SpanId": "1", "X-B3-Sampled": "1",
                                                      "Accept": "text/*;q=0.2, */*"}),
    client.build_request("OPTIONS", "/api/swagger.json"),
    client.build_request("GET", "/webjars/../../etc/passwd"),
]


def mutate(rng: random.Random, data: bytes) -> bytes:
    b = bytearray(data)
    for _ in range(rng.randint(1, 6)):
        op = rng.randint(0, 5)
        if op == 0 and b:
            b[rng.randrange(len(b))] = rng.randrange(256)
        elif op == 1 and b:
            del b[rng.randrange(len(b))]
        elif op == 2:
            b.insert(rng.randrange(len(b) + 1), rng.randrange(256))
        elif op == 3 and b:
            i = rng.randrange(len(b))
            b[i:i] = b[i:i + rng.randint(1, 40)] * rng.randint(1, 4)
        elif op == 4:
            i = rng.randrange(len(b) + 1)
            b[i:i] = rng.choice([b"\r\n", b"\n", b"\r", b":", b" ", b"\x00", b"%", b"HTTP/1.1", b"Content-Length: 99999999999999999999",
                                 b"Transfer-Encoding: chunked\r\n", b"\r\n\r\n"])
        elif op == 5 and len(b) > 4:
            del b[rng.randrange(len(b)):]
    return bytes(b)


def drain(s: socket.socket, budget: float = 0.3) -> None:
    """Reads whatever comes back, for a bounded time (a mutated request may have opened the
    never-ending metrics stream)."""
    import time
    deadline = time.time() + budget
    s.settimeout(0.05)
    try:
        while time.time() < deadline and s.recv(65536):
            pass
    except (socket.timeout, ConnectionError):
        pass


def alive(port: int) -> bool:
    r = client.get("127.0.0.1", port, "/api/health", timeout=5)
    return r.status == 200 and r.body == b"I'm ok"


def test_mutated_requests_never_take_the_server_down(hola_server):
    srv = hola_server(args=["--aloha-url", "http://127.0.0.1:9/", "--feign-max-attempts", "1", "--max-header-bytes", "8192"],
                      env={"ZIPKIN_SERVER_URL": "http://127.0.0.1:9"}, threads=2)
    rng = random.Random(20261106)
    for i in range(400):
        payload = mutate(rng, rng.choice(SEEDS))
        if rng.random() < 0.3:
            payload += mutate(rng, rng.choice(SEEDS))  # pipelined garbage
        s = socket.create_connection(("127.0.0.1", srv.port), timeout=2)
        try:
            # Sometimes dribble, sometimes blast.
            if rng.random() < 0.2:
                for k in range(0, len(payload), 7):
                    s.sendall(payload[k:k + 7])
            else:
                s.sendall(payload)
            if rng.random() < 0.5:
                s.shutdown(socket.SHUT_WR)
            drain(s)
        except OSError:  # the server may already have closed on us
            pass
        finally:
            s.close()
        if i % 50 == 49:
            assert alive(srv.port), "server stopped answering after %d fuzz cases" % (i + 1)
    assert alive(srv.port)
    assert srv.stop() == 0


def test_random_valid_pipelines_match_a_model(hola_server, fake_aloha):
    """Model-based check of the connection state machine: random sequences of *valid* requests,
    pipelined on one connection and cut into random TCP segments, must produce exactly the
    responses the same requests produce one at a time, in order. Includes deferred requests
    (the downstream call) and requests with bodies in the middle of the pipeline."""
    fake_aloha.delay = 0.002
    srv = hola_server(args=["--aloha-url", fake_aloha.url], threads=2)
    chained = ('["Hola de unknown",' + ",".join('"%s"' % p for p in fake_aloha.payload) + "]").encode()
    menu = [
        (client.build_request("GET", "/api/hola"), (200, b"Hola de unknown")),
        (client.build_request("GET", "/api/hola?x=%d" % 7), (200, b"Hola de unknown")),
        (client.build_request("GET", "/api/health"), (200, b"I'm ok")),
        (client.build_request("GET", "/api/%68ola;v=1"), (200, b"Hola de unknown")),   # canonicalised before routing
        (client.build_request("GET", "/api/../api/./health"), (200, b"I'm ok")),
        (client.build_request("GET", "/api/hola%zz"), None),                           # 400, and the connection is closed
        (client.build_request("HEAD", "/api/hola"), (200, b"")),
        (client.build_request("GET", "/api/hola-chaining"), (200, chained)),
        (client.build_request("POST", "/api/hola", body=b"GET /api/health HTTP/1.1\r\n\r\n"), (405, b"")),
        (b"PUT /api/health HTTP/1.1\r\nHost: x\r\nTransfer-Encoding: chunked\r\n\r\n3\r\nabc\r\n0\r\n\r\n", (405, b"")),
        (client.build_request("GET", "/api/none"), (404, None)),
        (client.build_request("GET", "/"), (200, None)),
        (client.build_request("GET", "/api/hola", headers={"Accept": "image/png"}), (406, None)),
        (client.build_request("OPTIONS", "/api/health"), (200, b"HEAD, GET, OPTIONS")),
    ]
    rng = random.Random(4242)
    for round_no in range(25):
        picks = [rng.choice(menu) for _ in range(rng.randint(1, 30))]
        # A malformed path ends the connection: everything before it is answered, then a 400.
        fatal = next((i for i, p in enumerate(picks) if p[1] is None), None)
        if fatal is not None:
            picks = picks[:fatal + 1]
        wire = b"".join(p[0] for p in picks)
        with client.Connection("127.0.0.1", srv.port, timeout=10) as c:
            pos = 0
            while pos < len(wire):
                step = rng.choice([1, 2, 3, 7, 20, 100, 1000, len(wire)])
                c.send_raw(wire[pos:pos + step])
                pos += step
            for i, (raw, expected) in enumerate(picks):
                if expected is None:
                    r = c.read_response()
                    assert r.status == 400 and c.closed_by_peer(), (round_no, i)
                    break
                status, body = expected
                r = c.read_response(head_only=raw.startswith(b"HEAD"))
                assert r.status == status, (round_no, i, raw[:40], r.status)
                if body is not None:
                    assert r.body == body, (round_no, i, raw[:40], r.body)
            # Nothing extra was produced, and the connection is still in sync.
            if fatal is None:
                assert c.request("GET", "/api/health").body == b"I'm ok"


def test_random_bytes_and_huge_lines(hola_server):
    srv = hola_server(threads=1, args=["--max-header-bytes", "4096"])
    rng = random.Random(7)
    blobs = [bytes(rng.randrange(256) for _ in range(rng.randint(1, 3000))) for _ in range(60)]
    blobs += [b"GET /" + b"a" * 100000 + b" HTTP/1.1\r\n\r\n", b"\r\n" * 5000, b"GET / HTTP/1.1\r\n" + b"X: y\r\n" * 5000 + b"\r\n",
              b"\x16\x03\x01\x02\x00\x01\x00\x01\xfc\x03\x03" + b"\x00" * 200]  # a TLS ClientHello on the plain port
    for blob in blobs:
        s = socket.create_connection(("127.0.0.1", srv.port), timeout=2)
        try:
            s.sendall(blob)
            drain(s)
        except OSError:
            pass
        finally:
            s.close()
    assert alive(srv.port)


def test_slow_reader_does_not_stall_others(hola_server):
    """A client that pipelines thousands of requests and never reads must not block the worker."""
    srv = hola_server(threads=1)
    hog = socket.create_connection(("127.0.0.1", srv.port))
    hog.setsockopt(socket.SOL_SOCKET, socket.SO_RCVBUF, 4096)
    hog.setblocking(False)
    sent = 0
    try:
        for _ in range(20000):
            sent += hog.send(GOOD * 50)
    except (BlockingIOError, ConnectionError):
        pass
    assert sent > 100000
    for _ in range(20):
        assert alive(srv.port)
    hog.close()
    assert alive(srv.port)
