"""ctypes bindings to ``build/libhola.so`` (C ABI in ``include/hola.h``).

Every wrapper fails loudly when the library is missing or a symbol is absent: there is no pure
Python fallback, so a test that passes has exercised the native code.
"""
from __future__ import annotations

import ctypes
import json
import struct
from ctypes import POINTER, byref, c_char_p, c_double, c_int, c_int64, c_size_t, c_uint32, c_uint64, c_void_p
from typing import Dict, List, Optional, Sequence, Tuple

from . import build, paths

_lib: Optional[ctypes.CDLL] = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        # HOLA_LIB_VARIANT=cov|asan|tsan loads the library of that build (tools/coverage.sh counts
        # the lines the C-ABI tests reach that way).
        import os
        variant = os.environ.get("HOLA_LIB_VARIANT", "release")
        p = paths(variant).libhola
        if not p.exists():
            build(variant)
        _lib = ctypes.CDLL(str(p))
        _declare(_lib)
    return _lib


def _declare(L: ctypes.CDLL) -> None:
    out = (POINTER(c_void_p), POINTER(c_size_t))
    L.hola_version.restype = c_char_p
    L.hola_free.argtypes = [c_void_p]
    L.hola_properties_parse.argtypes = [c_char_p, c_size_t, *out]
    L.hola_format.argtypes = [c_char_p, c_size_t, c_char_p, c_size_t, *out]
    L.hola_json_encode_list.argtypes = [c_char_p, c_size_t, *out]
    L.hola_json_decode_list.argtypes = [c_char_p, c_size_t, *out]
    L.hola_config_new.restype = c_void_p
    L.hola_config_free.argtypes = [c_void_p]
    L.hola_config_set_system_property.argtypes = [c_void_p, c_char_p, c_char_p]
    L.hola_config_add_properties_file.argtypes = [c_void_p, c_char_p]
    L.hola_config_resolve.argtypes = [c_void_p, c_char_p, *out]
    L.hola_config_greeting.argtypes = [c_void_p, c_char_p, c_int64, *out, POINTER(c_int)]
    L.hola_config_file_reads.argtypes = [c_void_p]
    L.hola_config_file_reads.restype = c_uint64
    L.hola_breaker_new.argtypes = [c_char_p]
    L.hola_breaker_new.restype = c_void_p
    L.hola_breaker_free.argtypes = [c_void_p]
    L.hola_breaker_admit.argtypes = [c_void_p, c_int64]
    L.hola_breaker_complete.argtypes = [c_void_p, c_int64, c_int, c_int]
    L.hola_breaker_is_open.argtypes = [c_void_p, c_int64]
    L.hola_breaker_health.argtypes = [c_void_p, c_int64, POINTER(c_int64), POINTER(c_int64), POINTER(c_int)]
    L.hola_breaker_command_json.argtypes = [c_void_p, c_int64, *out]
    L.hola_breaker_pool_json.argtypes = [c_void_p, c_int64, *out]
    L.hola_breaker_timeout_ms.argtypes = [c_void_p]
    L.hola_b3_server_decision.argtypes = [c_char_p, c_char_p, c_char_p, c_char_p, c_uint64, POINTER(c_uint64 * 3)]
    L.hola_span_json.argtypes = [c_uint64, c_uint64, c_uint64, c_int, c_int, c_int, c_int64, c_int64, c_int, c_char_p,
                                 c_char_p, c_uint32, *out]
    L.hola_http_parse_request.argtypes = [c_char_p, c_size_t, *out]
    L.hola_http_parse_response.argtypes = [c_char_p, c_size_t, *out]
    L.hola_http_dechunk.argtypes = [c_char_p, c_size_t, *out, POINTER(c_size_t)]
    L.hola_http_accept_allows.argtypes = [c_char_p, c_char_p, c_char_p]
    L.hola_http_date.argtypes = [c_int64, c_char_p]
    L.hola_histogram_new.restype = c_void_p
    L.hola_histogram_free.argtypes = [c_void_p]
    L.hola_histogram_record.argtypes = [c_void_p, c_uint64]
    L.hola_histogram_percentile.argtypes = [c_void_p, c_double]
    L.hola_histogram_percentile.restype = c_uint64
    L.hola_histogram_count.argtypes = [c_void_p]
    L.hola_histogram_count.restype = c_uint64
    L.hola_service_start.argtypes = [c_int, POINTER(c_char_p), POINTER(c_void_p)]
    L.hola_service_start.restype = c_void_p
    L.hola_service_port.argtypes = [c_void_p]
    L.hola_service_threads.argtypes = [c_void_p]
    L.hola_service_stats.argtypes = [c_void_p, POINTER(c_uint64), POINTER(c_uint64)]
    L.hola_service_trace_stats.argtypes = [c_void_p, POINTER(c_uint64), POINTER(c_uint64)]
    L.hola_service_stats_json.argtypes = [c_void_p, POINTER(c_void_p), POINTER(c_size_t)]
    L.hola_service_stop.argtypes = [c_void_p]


def _take(ptr: c_void_p, length: c_size_t) -> bytes:
    data = ctypes.string_at(ptr.value, length.value) if ptr.value else b""
    lib().hola_free(ptr)
    return data


def _call_out(fn, *args) -> Tuple[int, bytes]:
    ptr, length = c_void_p(), c_size_t()
    rc = fn(*args, byref(ptr), byref(length))
    return rc, _take(ptr, length)


def _as_bytes(s) -> bytes:
    return s if isinstance(s, bytes) else s.encode("utf-8")


def pack_list(items: Sequence[Optional[bytes]]) -> bytes:
    out = struct.pack("<I", len(items))
    for it in items:
        if it is None:
            out += b"\x00" + struct.pack("<I", 0)
        else:
            out += b"\x01" + struct.pack("<I", len(it)) + it
    return out


def unpack_list(data: bytes) -> List[Optional[bytes]]:
    (count,) = struct.unpack_from("<I", data, 0)
    pos = 4
    items: List[Optional[bytes]] = []
    for _ in range(count):
        present = data[pos]
        (n,) = struct.unpack_from("<I", data, pos + 1)
        pos += 5
        items.append(data[pos:pos + n] if present else None)
        pos += n
    return items


class NativeError(Exception):
    pass


def version() -> str:
    return lib().hola_version().decode()


def properties_parse(text) -> List[Tuple[bytes, bytes]]:
    """Entries of a java.util.Properties text in file order. Raises NativeError on malformed \\uXXXX."""
    data = _as_bytes(text)
    rc, out = _call_out(lib().hola_properties_parse, data, len(data))
    if rc != 0:
        raise NativeError(out.decode())
    flat = unpack_list(out)
    return [(flat[i], flat[i + 1]) for i in range(0, len(flat), 2)]


def java_format(pattern, arg) -> bytes:
    """String.format(pattern, arg) for one string argument. Raises NativeError with the JVM exception name."""
    p, a = _as_bytes(pattern), _as_bytes(arg)
    rc, out = _call_out(lib().hola_format, p, len(p), a, len(a))
    if rc != 0:
        raise NativeError(out.decode())
    return out


def json_encode_list(items: Sequence[Optional[bytes]]) -> bytes:
    packed = pack_list([None if i is None else _as_bytes(i) for i in items])
    rc, out = _call_out(lib().hola_json_encode_list, packed, len(packed))
    if rc != 0:
        raise NativeError(out.decode())
    return out


def json_decode_list(text) -> List[Optional[bytes]]:
    data = _as_bytes(text)
    rc, out = _call_out(lib().hola_json_decode_list, data, len(data))
    if rc != 0:
        raise NativeError(out.decode())
    return unpack_list(out)


class Config:
    """The service's configuration stack (core/config.h) as an object."""

    def __init__(self):
        self._h = lib().hola_config_new()

    def close(self):
        if self._h:
            lib().hola_config_free(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def set_system_property(self, key: str, value: str) -> None:
        lib().hola_config_set_system_property(self._h, key.encode(), value.encode())

    def add_properties_file(self, path: str) -> None:
        lib().hola_config_add_properties_file(self._h, path.encode())

    def resolve(self, key: str) -> Optional[bytes]:
        rc, out = _call_out(lib().hola_config_resolve, self._h, key.encode())
        return out if rc == 1 else None

    def greeting(self, hostname: str, now_ns: int) -> Tuple[bytes, bool]:
        ptr, length, refreshed = c_void_p(), c_size_t(), c_int()
        rc = lib().hola_config_greeting(self._h, hostname.encode(), now_ns, byref(ptr), byref(length), byref(refreshed))
        out = _take(ptr, length)
        if rc != 0:
            raise NativeError(out.decode())
        return out, bool(refreshed.value)

    @property
    def file_reads(self) -> int:
        return lib().hola_config_file_reads(self._h)


RUN, SHORT_CIRCUITED, POOL_REJECTED = 0, 1, 2
SUCCESS, FAILURE, TIMEOUT = 0, 1, 2


class Breaker:
    """The Hystrix-style command (core/breaker.h) driven by a caller-supplied clock."""

    def __init__(self, config_text: Optional[str] = None):
        self._h = lib().hola_breaker_new(config_text.encode() if config_text else None)

    def close(self):
        if self._h:
            lib().hola_breaker_free(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def admit(self, now_ms: int) -> int:
        return lib().hola_breaker_admit(self._h, now_ms)

    def complete(self, now_ms: int, outcome: int, execute_ms: int = 1) -> None:
        lib().hola_breaker_complete(self._h, now_ms, outcome, execute_ms)

    def is_open(self, now_ms: int) -> bool:
        return bool(lib().hola_breaker_is_open(self._h, now_ms))

    def health(self, now_ms: int) -> Tuple[int, int, int]:
        total, errors, pct = c_int64(), c_int64(), c_int()
        lib().hola_breaker_health(self._h, now_ms, byref(total), byref(errors), byref(pct))
        return total.value, errors.value, pct.value

    def command_json(self, now_ms: int) -> Dict:
        _, out = _call_out(lib().hola_breaker_command_json, self._h, now_ms)
        return json.loads(out)

    def pool_json(self, now_ms: int) -> Dict:
        _, out = _call_out(lib().hola_breaker_pool_json, self._h, now_ms)
        return json.loads(out)

    @property
    def timeout_ms(self) -> int:
        return lib().hola_breaker_timeout_ms(self._h)


def b3_server_decision(trace_id: Optional[str], span_id: Optional[str], parent_span_id: Optional[str],
                       sampled: Optional[str], seed: int = 1) -> Dict:
    enc = lambda s: None if s is None else s.encode()
    ids = (c_uint64 * 3)()
    flags = lib().hola_b3_server_decision(enc(trace_id), enc(span_id), enc(parent_span_id), enc(sampled), seed, byref(ids))
    return {"sampled": bool(flags & 1), "joined": bool(flags & 2), "has_parent": bool(flags & 4),
            "trace_id": ids[0], "span_id": ids[1], "parent_id": ids[2]}


def span_json(trace_id: int, span_id: int, parent_id: int = 0, has_parent: bool = False, client: bool = False,
              shared: bool = False, start_us: int = 0, end_us: int = 0, status: int = 200, url: str = "",
              service_name: str = "hola", ipv4: int = 0) -> Dict:
    _, out = _call_out(lib().hola_span_json, trace_id, span_id, parent_id, int(has_parent), int(client), int(shared),
                       start_us, end_us, status, url.encode(), service_name.encode(), ipv4)
    return json.loads(out)


def http_parse_request(data: bytes) -> Tuple[str, Optional[Dict]]:
    rc, out = _call_out(lib().hola_http_parse_request, data, len(data))
    if rc == 0:
        return "complete", json.loads(out)
    return ("incomplete" if rc == 1 else "bad"), None


def http_parse_response(data: bytes) -> Tuple[str, Optional[Dict]]:
    rc, out = _call_out(lib().hola_http_parse_response, data, len(data))
    if rc == 0:
        return "complete", json.loads(out)
    return ("incomplete" if rc == 1 else "bad"), None


def http_dechunk(data: bytes) -> Tuple[str, bytes, int]:
    ptr, length, consumed = c_void_p(), c_size_t(), c_size_t()
    rc = lib().hola_http_dechunk(data, len(data), byref(ptr), byref(length), byref(consumed))
    return {0: "done", 1: "incomplete", 2: "bad"}[rc], _take(ptr, length), consumed.value


def http_accept_allows(accept: str, type_: str, subtype: str) -> bool:
    return bool(lib().hola_http_accept_allows(accept.encode(), type_.encode(), subtype.encode()))


def http_date(unix_seconds: int) -> str:
    buf = ctypes.create_string_buffer(30)
    lib().hola_http_date(unix_seconds, buf)
    return buf.value.decode()


class Histogram:
    def __init__(self):
        self._h = lib().hola_histogram_new()

    def __del__(self):
        if self._h:
            lib().hola_histogram_free(self._h)
            self._h = None

    def record(self, value: int) -> None:
        lib().hola_histogram_record(self._h, value)

    def percentile(self, pct: float) -> int:
        return lib().hola_histogram_percentile(self._h, pct)

    @property
    def count(self) -> int:
        return lib().hola_histogram_count(self._h)


class Service:
    """The whole service running inside this process (include/hola.h: hola_service_*)."""

    def __init__(self, args: Sequence[str] = ("--port", "0", "--bind", "127.0.0.1", "--threads", "2", "--quiet")):
        self.args = [str(a) for a in args]
        self._h = None
        self.port = 0

    def start(self) -> "Service":
        argv = (c_char_p * len(self.args))(*[a.encode() for a in self.args])
        err = c_void_p()
        handle = lib().hola_service_start(len(self.args), argv, byref(err))
        if not handle:
            message = ctypes.string_at(err.value).decode() if err.value else "unknown error"
            lib().hola_free(err)
            raise NativeError(message)
        self._h = handle
        self.port = lib().hola_service_port(handle)
        return self

    def stop(self) -> None:
        if self._h:
            lib().hola_service_stop(self._h)
            self._h = None

    def stats(self) -> Dict[str, int]:
        req, conns = c_uint64(), c_uint64()
        lib().hola_service_stats(self._h, byref(req), byref(conns))
        return {"requests": req.value, "connections": conns.value}

    def trace_stats(self) -> Dict[str, int]:
        rep, drop = c_uint64(), c_uint64()
        lib().hola_service_trace_stats(self._h, byref(rep), byref(drop))
        return {"reported": rep.value, "dropped": drop.value}

    def stats_json(self) -> Dict:
        """Totals, CPU seconds, span counts and one entry per worker thread (see include/hola.h)."""
        import json
        _, data = _call_out(lib().hola_service_stats_json, self._h)
        return json.loads(data)

    def __enter__(self) -> "Service":
        return self.start()

    def __exit__(self, *exc) -> None:
        self.stop()
