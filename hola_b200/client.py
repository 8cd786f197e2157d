"""A deliberately small HTTP/1.1 client on raw sockets.

The tests assert on exact bytes (status line, header set, body) and exercise things ordinary
clients hide: pipelining, partial requests, half-closed connections, HTTP/1.0. SURVEY.md 2.2
pins the observable contract of every endpoint; this is the tool that checks it.
"""
from __future__ import annotations

import socket
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple


@dataclass
class Response:
    status: int
    reason: str
    version: str
    headers: List[Tuple[str, str]]
    body: bytes
    raw_head: bytes = b""

    def header(self, name: str) -> Optional[str]:
        for k, v in self.headers:
            if k.lower() == name.lower():
                return v
        return None

    def header_names(self) -> List[str]:
        return [k for k, _ in self.headers]

    @property
    def text(self) -> str:
        return self.body.decode("utf-8")


class Connection:
    """One TCP connection; requests are written verbatim, responses parsed strictly."""

    def __init__(self, host: str, port: int, timeout: float = 5.0):
        self.sock = socket.create_connection((host, port), timeout=timeout)
        self.sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        self.buf = b""

    def close(self) -> None:
        try:
            self.sock.close()
        except OSError:
            pass

    def __enter__(self) -> "Connection":
        return self

    def __exit__(self, *exc) -> None:
        self.close()

    def send_raw(self, data: bytes) -> None:
        self.sock.sendall(data)

    def request(self, method: str, target: str, headers: Optional[Dict[str, str]] = None, body: bytes = b"",
                version: str = "HTTP/1.1") -> Response:
        self.send_raw(build_request(method, target, headers, body, version))
        return self.read_response(head_only=(method == "HEAD"))

    def _fill(self) -> bool:
        chunk = self.sock.recv(65536)
        if not chunk:
            return False
        self.buf += chunk
        return True

    def read_response(self, head_only: bool = False) -> Response:
        while b"\r\n\r\n" not in self.buf:
            if not self._fill():
                raise ConnectionError("connection closed before a full response head (got %r)" % self.buf[:200])
        head, _, rest = self.buf.partition(b"\r\n\r\n")
        self.buf = rest
        lines = head.split(b"\r\n")
        version, _, tail = lines[0].partition(b" ")
        code, _, reason = tail.partition(b" ")
        headers: List[Tuple[str, str]] = []
        for line in lines[1:]:
            k, _, v = line.partition(b":")
            headers.append((k.decode("latin-1"), v.strip().decode("latin-1")))
        resp = Response(int(code), reason.decode("latin-1"), version.decode("latin-1"), headers, b"", head + b"\r\n\r\n")
        status = resp.status
        if head_only or status in (204, 304) or 100 <= status < 200:
            return resp
        te = (resp.header("Transfer-Encoding") or "").lower()
        if "chunked" in te:
            resp.body = self._read_chunked()
        elif resp.header("Content-Length") is not None:
            n = int(resp.header("Content-Length"))
            while len(self.buf) < n:
                if not self._fill():
                    raise ConnectionError("connection closed inside a body")
            resp.body, self.buf = self.buf[:n], self.buf[n:]
        else:
            while self._fill():
                pass
            resp.body, self.buf = self.buf, b""
        return resp

    def _read_line(self) -> bytes:
        while b"\r\n" not in self.buf:
            if not self._fill():
                raise ConnectionError("connection closed inside a chunk header")
        line, _, self.buf = self.buf.partition(b"\r\n")
        return line

    def read_chunk(self) -> Optional[bytes]:
        """Reads one chunk of a chunked body; ``None`` on the terminating chunk."""
        size = int(self._read_line().split(b";")[0], 16)
        if size == 0:
            self._read_line()
            return None
        while len(self.buf) < size + 2:
            if not self._fill():
                raise ConnectionError("connection closed inside a chunk")
        data, self.buf = self.buf[:size], self.buf[size + 2:]
        return data

    def _read_chunked(self) -> bytes:
        out = b""
        while True:
            chunk = self.read_chunk()
            if chunk is None:
                return out
            out += chunk

    def closed_by_peer(self, timeout: float = 2.0) -> bool:
        """True when the peer has closed (EOF) within ``timeout``; False when data or silence."""
        self.sock.settimeout(timeout)
        try:
            data = self.sock.recv(1)
        except socket.timeout:
            return False
        except OSError:
            return True
        if data:
            self.buf += data
            return False
        return True


def build_request(method: str, target: str, headers: Optional[Dict[str, str]] = None, body: bytes = b"",
                  version: str = "HTTP/1.1") -> bytes:
    hdrs = {"Host": "localhost"}
    if headers:
        for k, v in headers.items():
            # allow callers to override Host regardless of case
            for existing in list(hdrs):
                if existing.lower() == k.lower():
                    del hdrs[existing]
            hdrs[k] = v
    if body and not any(k.lower() in ("content-length", "transfer-encoding") for k in hdrs):
        hdrs["Content-Length"] = str(len(body))
    out = "%s %s %s\r\n" % (method, target, version)
    for k, v in hdrs.items():
        if v is not None:
            out += "%s: %s\r\n" % (k, v)
    return out.encode("latin-1") + b"\r\n" + body


def get(host: str, port: int, target: str, headers: Optional[Dict[str, str]] = None, method: str = "GET",
        timeout: float = 5.0) -> Response:
    with Connection(host, port, timeout) as c:
        return c.request(method, target, headers)
