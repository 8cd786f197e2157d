"""Fake neighbours of the hola service, for tests and for trying things out by hand.

* :class:`FakeAloha`  - the downstream `aloha` service (GET /api/aloha-chaining -> JSON list), with
  switchable misbehaviour: slow, error statuses, garbage, dropped connections, raw byte scripts.
* :class:`FakeZipkin` - a Zipkin collector that records what is POSTed to /api/v1/spans.
* :class:`FakeDns`    - a DNS server (UDP + TCP) answering A/AAAA queries from a dict.

``python -m hola_b200.fakes`` starts a fake aloha and a fake Zipkin and prints how to point a
hola instance at them.
"""
from __future__ import annotations

import json
import socket
import struct
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Dict, List


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


class FakeDns:
    """Answers A queries from a dict; everything else is NXDOMAIN. Optionally slow, silent, or
    truncating (TC bit over UDP, full answer over TCP on the same port)."""

    def __init__(self, records):
        self.records = {k.lower(): v for k, v in records.items()}
        self.delay = 0.0
        self.silent = False
        self.truncate_udp = False
        self.stray_first = None  # (name, ip): first send an answer to this *other* question under the query's id
        self.queries = []
        self.tcp_queries = []
        # UDP and TCP on the same port number; another socket may hold the TCP side of the port
        # the kernel picked for UDP, so try again until both can be had.
        for _ in range(50):
            self.sock = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
            self.sock.bind(("127.0.0.1", 0))
            self.port = self.sock.getsockname()[1]
            self.tcp = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
            self.tcp.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            try:
                self.tcp.bind(("127.0.0.1", self.port))
                break
            except OSError:
                self.sock.close()
                self.tcp.close()
        else:
            raise RuntimeError("no port free for both UDP and TCP")
        self.sock.settimeout(0.2)
        self.tcp.listen(16)
        self.tcp.settimeout(0.2)
        self._stop = False
        self.thread = threading.Thread(target=self._serve, daemon=True)
        self.thread.start()
        self.tcp_thread = threading.Thread(target=self._serve_tcp, daemon=True)
        self.tcp_thread.start()

    def _answer(self, data, truncated=False):
        qid = data[:2]
        pos, labels = 12, []
        while data[pos]:
            n = data[pos]
            labels.append(data[pos + 1:pos + 1 + n].decode())
            pos += 1 + n
        name = ".".join(labels).lower()
        question = data[12:pos + 5]
        (qtype,) = struct.unpack(">H", data[pos + 1:pos + 3])
        ip = self.records.get(name)
        if truncated:
            return name, qid + b"\x83\x80" + b"\x00\x01\x00\x00\x00\x00\x00\x00" + question
        if ip is None:
            return name, qid + b"\x81\x83" + b"\x00\x01\x00\x00\x00\x00\x00\x00" + question
        is_v6 = ":" in ip
        if (qtype == 1 and is_v6) or (qtype == 28 and not is_v6) or qtype not in (1, 28):
            # The name exists but has no record of the requested type: NOERROR, zero answers.
            return name, qid + b"\x81\x80" + b"\x00\x01\x00\x00\x00\x00\x00\x00" + question
        rdata = socket.inet_pton(socket.AF_INET6, ip) if is_v6 else socket.inet_aton(ip)
        answer = b"\xc0\x0c" + struct.pack(">HHIH", qtype, 1, 30, len(rdata)) + rdata
        return name, qid + b"\x81\x80" + b"\x00\x01\x00\x01\x00\x00\x00\x00" + question + answer

    def _serve(self):
        while not self._stop:
            try:
                data, addr = self.sock.recvfrom(2048)
            except socket.timeout:
                continue
            except OSError:
                return
            name, reply = self._answer(data, truncated=self.truncate_udp)
            self.queries.append(name)
            if self.silent:
                continue
            if self.delay:
                time.sleep(self.delay)
            if self.stray_first:
                # What an off-path spoofer who guessed the id would send: the right id, somebody
                # else's question. A resolver must ignore it and wait for the real answer.
                other, ip = self.stray_first
                qname = b"".join(bytes([len(l)]) + l.encode() for l in other.split(".")) + b"\x00"
                qtype = data[-4:-2]
                self.sock.sendto(data[:2] + b"\x81\x80" + b"\x00\x01\x00\x01\x00\x00\x00\x00" + qname + qtype + b"\x00\x01" +
                                 b"\xc0\x0c" + qtype + b"\x00\x01" + struct.pack(">IH", 30, 4) + socket.inet_aton(ip), addr)
                time.sleep(0.02)
            self.sock.sendto(reply, addr)

    def _serve_tcp(self):
        while not self._stop:
            try:
                conn, _ = self.tcp.accept()
            except socket.timeout:
                continue
            except OSError:
                return
            with conn:
                conn.settimeout(2)
                try:
                    head = conn.recv(2)
                    (n,) = struct.unpack(">H", head)
                    data = b""
                    while len(data) < n:
                        data += conn.recv(n - len(data))
                    name, reply = self._answer(data)
                    self.tcp_queries.append(name)
                    conn.sendall(struct.pack(">H", len(reply)) + reply)
                except (OSError, struct.error):
                    pass

    def stop(self):
        self._stop = True
        self.thread.join(timeout=2)
        self.tcp_thread.join(timeout=2)
        self.sock.close()
        self.tcp.close()


def main() -> None:
    aloha, zipkin = FakeAloha(), FakeZipkin()
    print("fake aloha  : %s   (GET /api/aloha-chaining -> %s)" % (aloha.url, json.dumps(aloha.payload)))
    print("fake zipkin : %s   (POST /api/v1/spans)" % zipkin.url)
    print()
    print("  ZIPKIN_SERVER_URL=%s ./build/hola --aloha-url %s --verbose" % (zipkin.url, aloha.url))
    print("  curl localhost:8080/api/hola-chaining")
    print()
    seen = 0
    try:
        while True:
            time.sleep(1)
            spans = zipkin.spans()
            for s in spans[seen:]:
                kind = s["annotations"][0]["value"]
                print("span %s trace=%s id=%s parent=%s %s" % ("server" if kind == "sr" else "client", s["traceId"], s["id"],
                                                             s.get("parentId", "-"), s["binaryAnnotations"][0]["value"]))
            seen = len(spans)
    except KeyboardInterrupt:
        pass
    finally:
        aloha.stop()
        zipkin.stop()


if __name__ == "__main__":
    main()
