"""HTTP/1.1 server behaviour that the reference inherits from Undertow and the rebuild has to
supply itself: keep-alive, pipelining, partial reads, request bodies, limits, shutdown."""
from __future__ import annotations

import socket
import threading
import time

import pytest

from hola_b200 import client

# The connection state machine has two I/O backends (src/net/server.h, IoBackend): every test here runs on both.
pytestmark = pytest.mark.usefixtures("both_io_backends")

HOLA = b"Hola de unknown"


def test_keep_alive_many_requests_one_connection(hola_server):
    srv = hola_server()
    with client.Connection("127.0.0.1", srv.port) as c:
        for i in range(200):
            r = c.request("GET", "/api/hola")
            assert r.status == 200 and r.body == HOLA
            assert r.header("Connection") == "keep-alive"


def test_pipelined_requests_answered_in_order(hola_server):
    srv = hola_server()
    with client.Connection("127.0.0.1", srv.port) as c:
        batch = b"".join(client.build_request("GET", t) for t in ("/api/hola", "/api/health", "/api/nothing", "/api/hola"))
        c.send_raw(batch * 10)
        for _ in range(10):
            assert c.read_response().body == HOLA
            assert c.read_response().body == b"I'm ok"
            assert c.read_response().status == 404
            assert c.read_response().body == HOLA


def test_request_split_into_single_bytes(hola_server):
    srv = hola_server()
    with client.Connection("127.0.0.1", srv.port) as c:
        for byte in client.build_request("GET", "/api/hola"):
            c.send_raw(bytes([byte]))
        assert c.read_response().body == HOLA
        # ... and the connection is still usable afterwards.
        assert c.request("GET", "/api/health").body == b"I'm ok"


def test_connection_close_and_http10(hola_server):
    srv = hola_server()
    with client.Connection("127.0.0.1", srv.port) as c:
        r = c.request("GET", "/api/hola", headers={"Connection": "close"})
        assert r.status == 200 and r.header("Connection") == "close"
        assert c.closed_by_peer()
    with client.Connection("127.0.0.1", srv.port) as c:
        r = c.request("GET", "/api/hola", version="HTTP/1.0")
        assert r.status == 200 and r.body == HOLA and r.header("Connection") == "close"
        assert c.closed_by_peer()
    with client.Connection("127.0.0.1", srv.port) as c:
        r = c.request("GET", "/api/hola", headers={"Connection": "keep-alive"}, version="HTTP/1.0")
        assert r.header("Connection") == "keep-alive"
        assert c.request("GET", "/api/health", headers={"Connection": "keep-alive"}, version="HTTP/1.0").body == b"I'm ok"


def test_request_bodies_are_skipped_not_misparsed(hola_server):
    srv = hola_server()
    with client.Connection("127.0.0.1", srv.port) as c:
        # A body that looks like a request must not be answered as one.
        r = c.request("POST", "/api/hola", body=b"GET /api/health HTTP/1.1\r\nHost: x\r\n\r\n")
        assert r.status == 405
        assert c.request("GET", "/api/hola").body == HOLA
        # Chunked body.
        c.send_raw(b"POST /api/hola HTTP/1.1\r\nHost: x\r\nTransfer-Encoding: chunked\r\n\r\n"
                   b"5\r\nhello\r\n1c\r\nGET /api/health HTTP/1.1\r\n\r\n\r\n0\r\n\r\n")
        assert c.read_response().status == 405
        assert c.request("GET", "/api/hola").body == HOLA
        # Body arriving in pieces after the head.
        c.send_raw(b"POST /api/health HTTP/1.1\r\nHost: x\r\nContent-Length: 10\r\n\r\n01234")
        assert c.read_response().status == 405
        c.send_raw(b"56789")
        assert c.request("GET", "/api/health").body == b"I'm ok"


def test_malformed_request_gets_400_and_close(hola_server):
    srv = hola_server()
    for raw in (b"NONSENSE\r\n\r\n", b"GET /api/hola HTTP/9.9\r\n\r\n", b"GET /api/hola HTTP/1.1\r\nBroken Header\r\n\r\n"):
        with client.Connection("127.0.0.1", srv.port) as c:
            c.send_raw(raw)
            r = c.read_response()
            assert r.status == 400 and r.header("Connection") == "close"
            assert c.closed_by_peer()


def test_oversized_head_gets_431(hola_server):
    srv = hola_server(args=["--max-header-bytes", "4096"])
    with client.Connection("127.0.0.1", srv.port) as c:
        c.send_raw(b"GET /api/hola HTTP/1.1\r\nHost: x\r\nX-Filler: " + b"a" * 8000)
        r = c.read_response()
        assert r.status == 431
        assert c.closed_by_peer()
    # A large-but-allowed head still works.
    with client.Connection("127.0.0.1", srv.port) as c:
        r = c.request("GET", "/api/hola", headers={"X-Filler": "b" * 3000})
        assert r.status == 200
    # With the default 64 KiB limit the receive buffer has to grow several times, also when the
    # head arrives in pieces, and the connection stays usable afterwards.
    srv2 = hola_server()
    with client.Connection("127.0.0.1", srv2.port) as c:
        big = client.build_request("GET", "/api/hola", headers={"X-Filler": "c" * 40000})
        for k in range(0, len(big), 9000):
            c.send_raw(big[k:k + 9000])
        assert c.read_response().body == HOLA
        assert c.request("GET", "/api/health").body == b"I'm ok"
        c.send_raw(client.build_request("GET", "/api/hola", headers={"X-Filler": "d" * 70000}))
        assert c.read_response().status == 431


def test_half_close_still_gets_the_answer(hola_server):
    srv = hola_server()
    with client.Connection("127.0.0.1", srv.port) as c:
        c.send_raw(client.build_request("GET", "/api/hola"))
        c.sock.shutdown(socket.SHUT_WR)
        assert c.read_response().body == HOLA
        assert c.closed_by_peer()


def test_many_concurrent_connections(hola_server):
    srv = hola_server(threads=4)
    errors = []

    def worker():
        try:
            with client.Connection("127.0.0.1", srv.port) as c:
                for _ in range(50):
                    if c.request("GET", "/api/hola").body != HOLA:
                        errors.append("bad body")
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=worker) for _ in range(32)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors


@pytest.mark.parametrize("mode", ["balanced", "reuseport", "rr"])
def test_accept_modes(hola_server, mode):
    srv = hola_server(args=["--accept-mode", mode], threads=3)
    conns = [client.Connection("127.0.0.1", srv.port) for _ in range(9)]
    try:
        for c in conns:
            assert c.request("GET", "/api/hola").body == HOLA
        stats = srv.stats()
        assert stats["accept_mode"] == mode and stats["connections_open"] == 9 and stats["requests"] == 9
        per_worker = [w["connections_open"] for w in stats["workers"]]
        assert sum(per_worker) == 9
        if mode != "reuseport":  # the kernel's hash makes no promise; the other two modes do
            assert per_worker == [3, 3, 3]
    finally:
        for c in conns:
            c.close()


def test_default_placement_spreads_keep_alive_connections_evenly(hola_server):
    """64 keep-alive clients over 16 workers is the benchmark's shape (BASELINE.json config 4 on a
    32-vCPU box). Undertow deals connections to its I/O threads in turn; hashing them over
    SO_REUSEPORT listeners leaves some workers with one connection and others with eight."""
    srv = hola_server(threads=16)
    conns = [client.Connection("127.0.0.1", srv.port) for _ in range(64)]
    try:
        for c in conns:
            assert c.request("GET", "/api/hola").body == HOLA
        stats = srv.stats()
        assert stats["accept_mode"] == "balanced"
        assert [w["connections_open"] for w in stats["workers"]] == [4] * 16
        assert [w["requests"] for w in stats["workers"]] == [4] * 16
        # Closing some and opening others keeps the spread within one connection.
        for c in conns[:20]:
            c.close()
        del conns[:20]
        time.sleep(0.2)
        conns += [client.Connection("127.0.0.1", srv.port) for _ in range(20)]
        for c in conns[-20:]:
            assert c.request("GET", "/api/health").body == b"I'm ok"
        per_worker = [w["connections_open"] for w in srv.stats()["workers"]]
        assert sum(per_worker) == 64 and max(per_worker) - min(per_worker) <= 1, per_worker
    finally:
        for c in conns:
            c.close()


def test_dual_stack_listener(hola_server):
    """`--bind ::` listens on IPv6 and IPv4 at once (the reference's container binds 0.0.0.0 only;
    IPv6-only pods need this)."""
    try:
        probe = socket.socket(socket.AF_INET6)
        probe.bind(("::1", 0))
        probe.close()
    except OSError:
        pytest.skip("no IPv6 loopback on this machine")
    srv = hola_server(args=["--bind", "::"])
    for family, host in ((socket.AF_INET6, "::1"), (socket.AF_INET, "127.0.0.1")):
        s = socket.socket(family, socket.SOCK_STREAM)
        s.settimeout(5)
        s.connect((host, srv.port))
        s.sendall(client.build_request("GET", "/api/hola", headers={"Connection": "close"}))
        data = b""
        while True:
            chunk = s.recv(4096)
            if not chunk:
                break
            data += chunk
        s.close()
        assert data.startswith(b"HTTP/1.1 200 OK") and data.endswith(HOLA), (host, data)


def test_idle_timeout(hola_server):
    srv = hola_server(args=["--idle-timeout", "300"])
    with client.Connection("127.0.0.1", srv.port) as c:
        assert c.request("GET", "/api/hola").status == 200
        assert c.closed_by_peer(timeout=3.0)


def test_graceful_shutdown_finishes_in_flight_and_exits_zero(hola_server):
    srv = hola_server()
    with client.Connection("127.0.0.1", srv.port) as c:
        assert c.request("GET", "/api/hola").status == 200
        t0 = time.time()
        code = srv.stop()
        assert code == 0
        assert time.time() - t0 < 3.0
        assert c.closed_by_peer()


def test_rolling_restart_loses_no_request(hola_server):
    """Zero-downtime replacement: a second instance joins the same port (SO_REUSEPORT), the first
    gets SIGTERM while 16 keep-alive clients hammer the port. The old instance answers what it has
    received, tells each client to reconnect (Connection: close) and exits 0; the clients land on
    the new instance; nobody sees a failed request."""
    import json
    import subprocess
    from hola_b200 import paths
    # A generous linger: on a starved machine a busy client can look idle for longer than the
    # default 100 ms, and a connection closed under a request in flight is the one thing that
    # would make this hand-over lossy.
    old = hola_server(threads=2, args=["--drain-linger-ms", "1000"])
    load = subprocess.Popen([str(paths().hola_load), "--json", "-c", "16", "-d", "3", "-w", "0", "-t", "2", "--expect-body",
                             "Hola de unknown", "http://127.0.0.1:%d/api/hola" % old.port], stdout=subprocess.PIPE, text=True)
    time.sleep(1.0)
    new = hola_server(args=["-Dswarm.http.port=%d" % old.port], threads=2)
    assert new.port == old.port
    time.sleep(0.3)
    assert old.stop() == 0
    out = load.communicate(timeout=30)[0]
    rep = json.loads(out.strip().splitlines()[-1])
    assert rep["errors"]["total"] == 0, rep["errors"]
    assert rep["reconnects"] >= 1  # clients really were moved over
    assert rep["requests"] > 1000
    # Throughput did not drop to zero in any interval.
    assert min(i["requests"] for i in rep["intervals"]) > 0


def test_date_header_is_current_and_well_formed(hola_server):
    import email.utils
    srv = hola_server()
    r = client.get("127.0.0.1", srv.port, "/api/hola")
    parsed = email.utils.parsedate_to_datetime(r.header("Date"))
    assert abs(parsed.timestamp() - time.time()) < 5
    assert r.header("Date").endswith(" GMT") and len(r.header("Date")) == 29


def test_descriptor_exhaustion_does_not_spin_or_die(hola_server, native_build):
    """With the fd limit reached, new connections wait in the backlog; the ones already open keep
    being served, the process does not burn a core on accept(), and service resumes afterwards."""
    import resource
    import subprocess
    import os

    # Launch with a tiny hard limit so the service cannot raise it.
    def limit():
        resource.setrlimit(resource.RLIMIT_NOFILE, (40, 40))

    env = {k: v for k, v in os.environ.items() if k not in ("HOSTNAME", "ZIPKIN_SERVER_URL")}
    ru_before = resource.getrusage(resource.RUSAGE_CHILDREN)
    proc = subprocess.Popen([str(native_build.hola), "--port", "0", "--bind", "127.0.0.1", "--threads", "1"],
                            stderr=subprocess.PIPE, text=True, env=env, preexec_fn=limit)
    try:
        line = proc.stderr.readline()
        port = int(line.split("127.0.0.1:")[1].split()[0])
        conns = []
        for _ in range(60):  # more than the limit allows
            try:
                c = client.Connection("127.0.0.1", port, timeout=1.0)
                conns.append(c)
            except OSError:
                break
        # The earliest connections still work.
        assert conns[0].request("GET", "/api/hola").body == HOLA
        assert conns[5].request("GET", "/api/health").body == b"I'm ok"
        time.sleep(0.5)
        # Free descriptors; the backlog drains and late connections get answered.
        for c in conns[:30]:
            c.close()
        time.sleep(0.3)
        late = conns[-1]
        late.sock.settimeout(3.0)
        assert late.request("GET", "/api/hola").body == HOLA
        for c in conns[30:]:
            c.close()
    finally:
        proc.terminate()
        out = proc.communicate(timeout=10)[1]
    assert proc.returncode == 0
    assert "pausing accepts" in out
    # RUSAGE_CHILDREN accumulates over every child this pytest process has reaped: take the delta.
    ru = resource.getrusage(resource.RUSAGE_CHILDREN)
    burnt = (ru.ru_utime + ru.ru_stime) - (ru_before.ru_utime + ru_before.ru_stime)
    assert burnt < 0.5, burnt  # a spinning accept loop would have burnt ~1 s of CPU


# -- regressions from the round-1 review ---------------------------------------------------------

def _cpu_seconds(pid: int) -> float:
    import os
    fields = open("/proc/%d/stat" % pid).read().rsplit(")", 1)[1].split()
    return (int(fields[11]) + int(fields[12])) / os.sysconf("SC_CLK_TCK")  # utime + stime


def test_stream_client_sending_garbage_does_not_pin_a_worker(hola_server):
    """A metrics-stream client that keeps writing used to leave the socket readable for ever:
    level-triggered epoll then spun one worker at 100 % CPU. The input is now read and dropped."""
    srv = hola_server(threads=1)
    c = client.Connection("127.0.0.1", srv.port)
    c.send_raw(client.build_request("GET", "/hystrix.stream?delay=100"))
    while b"\r\n\r\n" not in c.buf:
        assert c._fill()
    c.send_raw(b"x" * 80 * 1024)
    time.sleep(0.3)
    before = _cpu_seconds(srv.proc.pid)
    time.sleep(1.5)
    used = _cpu_seconds(srv.proc.pid) - before
    assert used < 0.3, "worker burned %.2f s of CPU in 1.5 s" % used
    # The stream is still alive and the worker still serves others.
    assert client.get("127.0.0.1", srv.port, "/api/hola").body == HOLA
    c.sock.settimeout(2)
    assert c.sock.recv(65536)
    c.close()


def test_deferred_request_with_full_buffer_and_half_closed_peer_does_not_spin(hola_server, fake_aloha):
    """While the downstream call of /api/hola-chaining is in flight the connection stops parsing.
    A client that pipelines more than the buffer holds and then half-closes used to leave
    EPOLLRDHUP armed on a socket whose EOF could not be read: a busy loop until the call settled.
    Every pipelined request must still be answered, in order."""
    fake_aloha.delay = 1.5
    srv = hola_server(args=["--aloha-url", fake_aloha.url, "--max-header-bytes", "1024",
                            "-Dhystrix.command.default.execution.isolation.thread.timeoutInMilliseconds=5000"], threads=1)
    c = client.Connection("127.0.0.1", srv.port, timeout=10)
    extra = 400  # ~16 KB of pipelined requests against a 5 KB buffer
    c.send_raw(client.build_request("GET", "/api/hola-chaining") + client.build_request("GET", "/api/hola") * extra)
    c.sock.shutdown(socket.SHUT_WR)
    time.sleep(0.3)
    before = _cpu_seconds(srv.proc.pid)
    time.sleep(0.9)
    used = _cpu_seconds(srv.proc.pid) - before
    assert used < 0.2, "worker burned %.2f s of CPU in 0.9 s while the call was pending" % used
    first = c.read_response()
    assert first.status == 200 and first.body.startswith(b'["Hola de unknown","Aloha mai')
    for _ in range(extra):
        assert c.read_response().body == HOLA
    assert c.closed_by_peer()
    c.close()


def test_half_close_behind_a_deferred_request_keeps_the_pipelined_requests(hola_server, fake_aloha):
    """A client writes a batch - the first request needs the downstream call, the rest are queued
    behind it - and half-closes straight away. The FIN reaches the server while the first request
    is still pending; the queued requests were sent before it and must all be answered."""
    fake_aloha.delay = 0.3
    srv = hola_server(args=["--aloha-url", fake_aloha.url], threads=1)
    c = client.Connection("127.0.0.1", srv.port, timeout=10)
    c.send_raw(client.build_request("GET", "/api/hola-chaining") + client.build_request("GET", "/api/hola") * 3 +
               client.build_request("GET", "/api/health"))
    c.sock.shutdown(socket.SHUT_WR)
    assert c.read_response().body.startswith(b'["Hola de unknown","Aloha mai')
    for _ in range(3):
        assert c.read_response().body == HOLA
    assert c.read_response().body == b"I'm ok"
    assert c.closed_by_peer()
    c.close()
    # ... and the plain case: a batch and a FIN with nothing deferred.
    c = client.Connection("127.0.0.1", srv.port, timeout=10)
    c.send_raw(client.build_request("GET", "/api/hola") * 5)
    c.sock.shutdown(socket.SHUT_WR)
    for _ in range(5):
        assert c.read_response().body == HOLA
    assert c.closed_by_peer()
    c.close()


def test_connection_slot_is_not_reused_within_one_batch_of_events(hola_server, fake_aloha):
    """One worker, stopped with SIGSTOP while three things queue up in this order: the downstream
    reply for deferred connection A, a new client B, and a reset of A. On resume the reply cannot
    be written to A (its slot is freed), B is accepted, and A's reset is delivered last. B must be
    answered: the freed slot may not be handed to B while A's stale event is still in the batch."""
    import os
    import signal
    import struct
    fake_aloha.delay = 0.4
    for _ in range(3):  # the order of readiness is the kernel's; repeat to be sure the window was hit
        srv = hola_server(args=["--aloha-url", fake_aloha.url], threads=1)
        a = client.Connection("127.0.0.1", srv.port)
        a.send_raw(client.build_request("GET", "/api/hola-chaining"))
        time.sleep(0.15)  # the downstream call is in flight
        os.kill(srv.proc.pid, signal.SIGSTOP)
        try:
            time.sleep(0.5)  # the downstream reply has arrived at the stopped process
            b = client.Connection("127.0.0.1", srv.port)
            b.send_raw(client.build_request("GET", "/api/hola"))
            time.sleep(0.05)
            a.sock.setsockopt(socket.SOL_SOCKET, socket.SO_LINGER, struct.pack("ii", 1, 0))
            a.sock.close()  # RST
            time.sleep(0.05)
        finally:
            os.kill(srv.proc.pid, signal.SIGCONT)
        r = b.read_response()
        assert r.status == 200 and r.body == HOLA
        assert b.request("GET", "/api/health").body == b"I'm ok"
        b.close()
        srv.stop()
