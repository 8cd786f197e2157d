"""Building blocks exercised through the C ABI (include/hola.h), with property-based tests.

Oracles are independent Python restatements of the specifications the reference relies on
(java.util.Properties#load, JSON via the standard library), not the C++ code under test.
"""
from __future__ import annotations

import json
import re
import string

import pytest
from hypothesis import given, settings, strategies as st

from hola_b200 import native


# ---- java.util.Properties: an independent, deliberately naive Python model ----------------------

def model_properties(text: str):
    """Straight transcription of the load(Reader) rules; quadratic and simple on purpose."""
    # 1. natural lines
    lines = re.split(r"\r\n|\r|\n", text)
    logical, i = [], 0
    while i < len(lines):
        line = lines[i].lstrip(" \t\f")
        i += 1
        if line == "" or line[0] in "#!":
            continue
        while True:
            m = re.search(r"\\+$", line)
            if m and len(m.group(0)) % 2 == 1:
                line = line[:-1]
                if i < len(lines):
                    line += lines[i].lstrip(" \t\f")
                    i += 1
                    continue
            break
        logical.append(line)
    out = []
    for line in logical:
        k = 0
        while k < len(line):
            ch = line[k]
            if ch == "\\":
                k += 2
                continue
            if ch in "=: \t\f":
                break
            k += 1
        key_raw = line[:min(k, len(line))]
        rest = line[min(k, len(line)):]
        rest = rest.lstrip(" \t\f")
        if rest[:1] in ("=", ":"):
            rest = rest[1:].lstrip(" \t\f")
        out.append((model_unescape(key_raw), model_unescape(rest)))
    return out


def model_unescape(s: str) -> str:
    out, i = [], 0
    while i < len(s):
        ch = s[i]
        if ch != "\\":
            out.append(ch)
            i += 1
            continue
        i += 1
        if i >= len(s):
            break
        e = s[i]
        i += 1
        if e == "u":
            hexs = s[i:i + 4]
            if len(hexs) < 4 or any(c not in string.hexdigits for c in hexs):
                raise ValueError("Malformed \\uxxxx encoding.")
            out.append(chr(int(hexs, 16)))
            i += 4
        else:
            out.append({"t": "\t", "n": "\n", "r": "\r", "f": "\f"}.get(e, e))
    return "".join(out)


def encode_like_jvm(s: str) -> bytes:
    # UTF-16 code units -> UTF-8, pairing surrogates, '?' for lone ones.
    return s.encode("utf-16", "surrogatepass").decode("utf-16", "replace").replace("�", "?").encode("utf-8")


_props_alphabet = st.sampled_from(list("abk =:#!\\\t\n\r\fuU0041eé€%s"))


@settings(max_examples=400, deadline=None)
@given(st.text(alphabet=_props_alphabet, max_size=60))
def test_properties_match_the_model(text):
    try:
        expected = model_properties(text)
    except ValueError:
        with pytest.raises(native.NativeError):
            native.properties_parse(text)
        return
    got = native.properties_parse(text)
    assert got == [(encode_like_jvm(k), encode_like_jvm(v)) for k, v in expected]


def test_properties_golden_vectors():
    assert native.properties_parse(b"hello=Hola de %s mediante el archivo de configuraciones\\n") == [
        (b"hello", b"Hola de %s mediante el archivo de configuraciones\n")]
    assert native.properties_parse("# c\n! c\n\nkey\\ with\\ spaces : value \\\n   continued\n") == [
        (b"key with spaces", b"value continued")]
    assert native.properties_parse("a\\u0020b=\\uD83D\\uDE00") == [(b"a b", "😀".encode())]
    with pytest.raises(native.NativeError, match="Malformed"):
        native.properties_parse("x=\\u12")


# ---- String.format ----------------------------------------------------------------------------

@settings(max_examples=300, deadline=None)
@given(st.text(alphabet=st.characters(blacklist_characters="%", blacklist_categories=("Cs",)), max_size=40),
       st.text(alphabet=st.characters(blacklist_categories=("Cs",)), max_size=20))
def test_format_without_directives_is_identity_and_s_substitutes(literal, host):
    assert native.java_format(literal, host) == literal.encode()
    assert native.java_format(literal + "%s" + literal, host) == (literal + host + literal).encode()
    assert native.java_format("%%" + literal, host) == ("%" + literal).encode()


@settings(max_examples=200, deadline=None)
@given(st.integers(min_value=1, max_value=40), st.text(alphabet=string.ascii_letters + "éß€", max_size=30), st.booleans())
def test_format_width_pads_like_python(width, host, left):
    spec = "%" + ("-" if left else "") + str(width) + "s"
    expected = host.ljust(width) if left else host.rjust(width)
    assert native.java_format(spec, host) == expected.encode()


@settings(max_examples=300, deadline=None)
@given(st.booleans(), st.one_of(st.none(), st.integers(min_value=1, max_value=30)), st.one_of(st.none(), st.integers(min_value=0, max_value=30)),
       st.text(alphabet=string.ascii_letters + string.digits + " éßñ€中", max_size=25), st.sampled_from(["s", "S"]))
def test_format_width_and_precision_agree_with_python(left, width, precision, host, conv):
    """For BMP text, Java's %[-][w][.p]s and Python's are the same function (width and precision
    count characters; precision truncates before padding). %S additionally upper-cases."""
    if left and width is None:
        return  # '-' without a width is an error on the JVM (covered elsewhere)
    spec = "%" + ("-" if left else "") + ("" if width is None else str(width)) + ("" if precision is None else "." + str(precision))
    expected = (spec + "s") % host
    if conv == "S":
        # str.upper() and String.toUpperCase() agree on this alphabet (including ß -> SS); upper-casing
        # happens after truncation, before padding.
        cut = host if precision is None else host[:precision]
        up = cut.upper()
        expected = up.ljust(width) if (left and width) else (up.rjust(width) if width else up)
    assert native.java_format("[" + spec + conv + "]", host) == ("[" + expected + "]").encode()


@pytest.mark.parametrize("pattern,exc", [
    ("%s %s", "MissingFormatArgumentException"),
    ("%d", "IllegalFormatConversionException"),
    ("%", "UnknownFormatConversionException"),
    ("%y", "UnknownFormatConversionException"),
    ("%-s", "MissingFormatWidthException"),
    ("%#s", "FormatFlagsConversionMismatchException"),
    ("%,s", "FormatFlagsConversionMismatchException"),
    ("%1$-#s", "MissingFormatWidthException"),
])
def test_format_rejections(pattern, exc):
    with pytest.raises(native.NativeError, match=exc):
        native.java_format(pattern, "h")


@settings(max_examples=300, deadline=None)
@given(st.text(alphabet=st.sampled_from(list("%sdnb-#+ 0,(<.$123tTYx%hH")), max_size=12))
def test_format_never_crashes_and_is_total(pattern):
    try:
        out = native.java_format(pattern, "host")
        assert isinstance(out, bytes)
    except native.NativeError as e:
        assert "Exception" in str(e)


# ---- JSON list codec ---------------------------------------------------------------------------

_json_text = st.text(alphabet=st.characters(blacklist_categories=("Cs",)), max_size=30)


@settings(max_examples=300, deadline=None)
@given(st.lists(st.one_of(st.none(), _json_text), max_size=6))
def test_json_encode_is_valid_and_compact(items):
    encoded = native.json_encode_list([None if i is None else i.encode() for i in items])
    assert json.loads(encoded) == items
    if all(i is None or all(ord(ch) < 0x7F and ch.isprintable() and ch not in '"\\' for ch in i) for i in items):
        assert encoded == json.dumps(items, separators=(",", ":")).encode()


@settings(max_examples=300, deadline=None)
@given(st.lists(_json_text, max_size=6), st.booleans(), st.integers(min_value=0, max_value=4))
def test_json_decode_accepts_what_python_writes(items, ascii_only, indent):
    text = json.dumps(items, ensure_ascii=ascii_only, indent=indent or None)
    assert native.json_decode_list(text) == [i.encode() for i in items]


@settings(max_examples=200, deadline=None)
@given(st.binary(max_size=40))
def test_json_decode_never_crashes(blob):
    try:
        native.json_decode_list(blob)
    except native.NativeError:
        pass


def test_json_scalar_coercion_like_jackson():
    assert native.json_decode_list('[1, 2.50, true, null, "x"]') == [b"1", b"2.50", b"true", None, b"x"]
    for bad in ('{"a":1}', '["a",', '[["x"]]', "nope", ""):
        with pytest.raises(native.NativeError):
            native.json_decode_list(bad)


# ---- configuration stack ------------------------------------------------------------------------

def test_config_greeting_cache_and_reload(tmp_path, monkeypatch):
    monkeypatch.delenv("hello", raising=False)
    conf = tmp_path / "t.properties"
    conf.write_text("hello=Uno %s\n")
    cfg = native.Config()
    assert cfg.greeting("h", 0) == (b"Hola de h", True)  # no conf yet
    cfg.set_system_property("conf", str(conf))
    s = 10 ** 9
    assert cfg.greeting("h", 1 * s) == (b"Hola de h", False)  # cached for 5 s
    assert cfg.greeting("h", 6 * s) == (b"Uno h", True)
    reads = cfg.file_reads
    for t in range(7, 11):
        assert cfg.greeting("h", t * s) == (b"Uno h", False)
    assert cfg.file_reads == reads  # the cache really prevents file I/O
    conf.write_text("hello=Dos %s\n")
    assert cfg.greeting("h", 12 * s) == (b"Dos h", True)
    monkeypatch.setenv("hello", "Env %s")
    assert cfg.greeting("h", 18 * s) == (b"Env h", True)
    cfg.set_system_property("hello", "Prop %s y %s")
    with pytest.raises(native.NativeError, match="MissingFormatArgument"):
        cfg.greeting("h", 24 * s)


def test_config_variable_expansion_and_static_file(tmp_path, monkeypatch):
    monkeypatch.delenv("hello", raising=False)
    extra = tmp_path / "apache-deltaspike.properties"
    extra.write_text("hello=Hola ${planet}\nplanet=Marte\n")
    cfg = native.Config()
    cfg.add_properties_file(str(extra))
    assert cfg.resolve("hello") == b"Hola Marte"
    cfg.set_system_property("planet", "Tierra")
    assert cfg.resolve("hello") == b"Hola Tierra"
    assert cfg.resolve("absent") is None


# ---- breaker through the ABI (state-machine properties) ------------------------------------------

@settings(max_examples=100, deadline=None)
@given(st.lists(st.tuples(st.integers(min_value=0, max_value=700), st.sampled_from([native.SUCCESS, native.FAILURE, native.TIMEOUT])),
                min_size=1, max_size=60))
def test_breaker_invariants(steps):
    b = native.Breaker()
    now = 1000
    admitted = short = 0
    for dt, outcome in steps:
        now += dt
        was_open = b.is_open(now)
        verdict = b.admit(now)
        if verdict == native.RUN:
            admitted += 1
            b.complete(now, outcome, 1)
            if outcome == native.SUCCESS:
                assert not b.is_open(now)  # a success always leaves the circuit closed
        else:
            assert verdict == native.SHORT_CIRCUITED and was_open
            short += 1
        total, errors, pct = b.health(now)
        assert 0 <= errors <= total <= admitted and 0 <= pct <= 100
    cmd = b.command_json(now + 20000)
    assert cmd["rollingCountShortCircuited"] == 0  # everything has aged out of the 10 s window


class BreakerModel:
    """Independent Python restatement of the breaker rules (Hystrix 1.5.1, reference settings):
    500 ms health buckets over a 10 s window, only closed buckets visible; open at >= 5 requests
    and >= 50 % errors; one trial per 5 s sleep window; success closes and wipes the window;
    bulkhead of 10 concurrent executions, rejections count as errors."""

    BUCKET, WINDOW, THRESHOLD, ERR_PCT, SLEEP, POOL = 500, 10000, 5, 50, 5000, 10

    def __init__(self):
        self.events = []  # (time, is_error) of everything that counts as a request
        self.origin = None  # buckets are aligned to the first event
        self.window_start = None  # after a reset, older events are forgotten
        self.open = False
        self.tested = 0
        self.active = 0

    def _visible(self, now):
        if self.origin is None:
            return []
        current_bucket_start = self.origin + ((now - self.origin) // self.BUCKET) * self.BUCKET
        oldest = current_bucket_start - self.WINDOW
        lo = oldest if self.window_start is None else max(oldest, self.window_start)
        return [e for (t, e) in self.events if lo <= t < current_bucket_start]

    def _touch(self, now):
        if self.origin is None:
            self.origin = now

    def is_open(self, now):
        self._touch(now)
        if self.open:
            return True
        seen = self._visible(now)
        total, errors = len(seen), sum(seen)
        if total < self.THRESHOLD or int(errors / total * 100) < self.ERR_PCT:
            return False
        self.open, self.tested = True, now
        return True

    def admit(self, now):
        self._touch(now)
        if self.is_open(now):
            if now > self.tested + self.SLEEP:
                self.tested = now
            else:
                return native.SHORT_CIRCUITED
        if self.active >= self.POOL:
            self.events.append((now, 1))
            return native.POOL_REJECTED
        self.active += 1
        return native.RUN

    def complete(self, now, outcome):
        self._touch(now)
        self.active -= 1
        self.events.append((now, 0 if outcome == native.SUCCESS else 1))
        if outcome == native.SUCCESS and self.open:
            self.open = False
            # The window restarts at `now`: bucket alignment moves with it.
            self.origin = now
            self.window_start = now
            self.events = []


@settings(max_examples=200, deadline=None)
@given(st.lists(st.tuples(st.integers(min_value=0, max_value=1500),      # time step
                          st.sampled_from(["call", "call", "call", "finish"]),
                          st.sampled_from([native.SUCCESS, native.FAILURE, native.FAILURE, native.TIMEOUT]),
                          st.integers(min_value=0, max_value=3)),        # hold the call open for this many later steps
                min_size=1, max_size=80))
def test_breaker_matches_independent_model(steps):
    real, model = native.Breaker(), BreakerModel()
    now = 10_000
    in_flight = []  # (remaining_steps, outcome)
    for dt, action, outcome, hold in steps:
        now += dt
        # Finish calls whose time has come (both sides see the same completions at the same time).
        still = []
        for remaining, out in in_flight:
            if remaining <= 0 or action == "finish":
                real.complete(now, out, 1)
                model.complete(now, out)
            else:
                still.append((remaining - 1, out))
        in_flight = still
        if action == "call":
            a, b = real.admit(now), model.admit(now)
            assert a == b, (now, a, b)
            if a == native.RUN:
                if hold == 0:
                    real.complete(now, outcome, 1)
                    model.complete(now, outcome)
                else:
                    in_flight.append((hold, outcome))
        assert real.is_open(now) == model.is_open(now), now


def test_breaker_custom_config_text():
    b = native.Breaker("hystrix.command.default.circuitBreaker.requestVolumeThreshold=2\n"
                       "hystrix.command.default.execution.isolation.thread.timeoutInMilliseconds=250\n")
    assert b.timeout_ms == 250
    for t in (1000, 1001):
        assert b.admit(t) == native.RUN
        b.complete(t, native.FAILURE)
    assert b.is_open(1600)


# ---- tracing, http, histogram ----------------------------------------------------------------------

def test_b3_decisions():
    d = native.b3_server_decision(None, None, None, None)
    assert d["sampled"] and not d["joined"] and d["trace_id"] == d["span_id"] != 0
    d = native.b3_server_decision("463ac35c9f6413ad", "a2fb4a1d1a96d312", None, "1")
    assert d["joined"] and d["trace_id"] == 0x463ac35c9f6413ad and not d["has_parent"]
    d = native.b3_server_decision("463ac35c9f6413ad", "a2fb4a1d1a96d312", None, None)
    assert not d["joined"]  # Brave 3.9 ignores ids when X-B3-Sampled is absent
    assert not native.b3_server_decision("1", "2", None, "0")["sampled"]
    assert not native.b3_server_decision("1", "2", None, "False")["sampled"]
    d = native.b3_server_decision("zz", "2", None, "1")
    assert d["sampled"] and not d["joined"]  # unparsable ids: start a new trace


def test_span_json_is_zipkin_v1():
    s = native.span_json(0xabc, 0xdef, 0x123, has_parent=True, client=True, start_us=10, end_us=35, status=503,
                         url="http://aloha:8080/api/aloha-chaining", ipv4=0xC0A80001)
    assert s["traceId"] == "0000000000000abc" and s["id"] == "0000000000000def" and s["parentId"] == "0000000000000123"
    assert s["timestamp"] == 10 and s["duration"] == 25
    assert [a["value"] for a in s["annotations"]] == ["cs", "cr"]
    assert s["annotations"][0]["endpoint"] == {"serviceName": "hola", "ipv4": "192.168.0.1"}
    assert {t["key"]: t["value"] for t in s["binaryAnnotations"]} == {
        "http.url": "http://aloha:8080/api/aloha-chaining", "http.status_code": "503"}


@settings(max_examples=200, deadline=None)
@given(st.binary(max_size=200))
def test_http_request_parser_never_crashes(blob):
    state, info = native.http_parse_request(blob)
    assert state in ("complete", "incomplete", "bad")
    if state == "complete":
        assert 0 < info["head_len"] <= len(blob)


@settings(max_examples=200, deadline=None)
@given(st.sampled_from(["GET", "HEAD", "POST", "OPTIONS", "DELETE"]),
       st.text(alphabet=string.ascii_letters + string.digits + "/-_.~", min_size=0, max_size=30),
       st.text(alphabet=string.ascii_letters + string.digits + "=&", max_size=20),
       st.lists(st.tuples(st.text(alphabet=string.ascii_letters + "-", min_size=1, max_size=12),
                          st.text(alphabet=string.ascii_letters + string.digits + " ,;=/*.", max_size=30)), max_size=6),
       st.binary(max_size=20))
def test_http_request_parser_roundtrip(method, path, query, headers, trailing):
    target = "/" + path + ("?" + query if query else "")
    head = "%s %s HTTP/1.1\r\n" % (method, target)
    head += "".join("%s: %s\r\n" % (k, v) for k, v in headers if k.lower() not in ("content-length", "transfer-encoding"))
    head += "\r\n"
    raw = head.encode() + trailing
    state, info = native.http_parse_request(raw)
    assert state == "complete"
    assert info["method"] == method and info["path"] == "/" + path and info["query"] == query
    assert info["head_len"] == len(head)
    # Any strict prefix of the head is incomplete, never an error and never a false "complete".
    for cut in range(0, len(head), 7):
        assert native.http_parse_request(raw[:cut])[0] == "incomplete"


def _model_canonical(raw: str):
    """Independent restatement of core/http.h canonicalise_path: cut matrix parameters, decode
    (not %2F/%5C), resolve dot and empty segments; None for a malformed escape or a decoded NUL."""
    import re
    trailing = len(raw) > 1 and raw.endswith("/")
    stack = []
    segs = raw[1:].split("/")
    for idx, seg in enumerate(segs):
        seg = seg.split(";", 1)[0]
        if re.search(r"%(?![0-9a-fA-F]{2})", seg):
            return None
        if re.search(r"%00", seg):
            return None
        seg = re.sub(r"%([0-9a-fA-F]{2})", lambda m: m.group(0) if m.group(1).lower() in ("2f", "5c") else chr(int(m.group(1), 16)), seg)
        last = idx == len(segs) - 1
        if seg == "" or seg == ".":
            if last and seg == ".":
                trailing = True
        elif seg == "..":
            if stack:
                stack.pop()
            if last:
                trailing = True
        else:
            stack.append(seg)
    out = "".join("/" + x for x in stack)
    if out == "" or (trailing and not out.endswith("/")):
        out += "/"
    return out


_PATH_PIECES = st.sampled_from(["api", "hola", "health", ".", "..", "", "%68ola", "hol%61", "%2e", "%2E%2e", ";a=1", "hola;v=2;w",
                                "..;x", "%2f", "%5C", "%3B", "a%20b", "~", "%7e", "%", "%4", "%zz", "%00", "x%41%42", "-_.!"])


@settings(max_examples=400, deadline=None)
@given(st.lists(_PATH_PIECES, min_size=0, max_size=7), st.booleans())
def test_canonical_path_matches_the_model(pieces, trailing_slash):
    raw = "/" + "/".join(pieces) + ("/" if trailing_slash else "")
    state, info = native.http_parse_request(("GET %s HTTP/1.1\r\n\r\n" % raw).encode("latin-1"))
    assert state == "complete" and info["path"] == raw
    want = _model_canonical(raw)
    got = info["canonical_path"]
    assert (got.encode("latin-1").decode("latin-1") if got is not None else None) == want, raw
    if want is not None:
        assert want.startswith("/") and "//" not in want
        assert not any(seg in (".", "..") for seg in want.split("/"))
    if info["path_is_canonical"]:
        assert want == raw  # the fast path is only taken when there is nothing to do


@settings(max_examples=200, deadline=None)
@given(st.lists(st.binary(min_size=1, max_size=50), max_size=6), st.binary(max_size=10))
def test_dechunk_roundtrip(chunks, rest):
    wire = b"".join(b"%x\r\n" % len(c) + c + b"\r\n" for c in chunks) + b"0\r\n\r\n"
    state, body, used = native.http_dechunk(wire + rest)
    assert (state, body, used) == ("done", b"".join(chunks), len(wire))
    if len(wire) > 1:
        assert native.http_dechunk(wire[:-1])[0] == "incomplete"


def test_http_date_and_accept():
    assert native.http_date(784111777) == "Sun, 06 Nov 1994 08:49:37 GMT"
    assert native.http_accept_allows("text/*;q=0.3, application/json", "text", "plain")
    assert not native.http_accept_allows("image/png", "text", "plain")


@settings(max_examples=50, deadline=None)
@given(st.lists(st.integers(min_value=1, max_value=10 ** 10), min_size=1, max_size=300))
def test_histogram_percentiles_within_one_percent(values):
    h = native.Histogram()
    for v in values:
        h.record(v)
    assert h.count == len(values)
    ordered = sorted(values)
    for pct in (50, 90, 99, 100):
        rank = max(1, min(len(ordered), int(pct / 100 * len(ordered) + 0.5)))
        exact = ordered[rank - 1]
        got = h.percentile(pct)
        assert exact <= got <= exact * 1.01 + 1, (pct, exact, got)


def test_service_in_process(monkeypatch):
    monkeypatch.delenv("ZIPKIN_SERVER_URL", raising=False)
    monkeypatch.delenv("HOSTNAME", raising=False)
    from hola_b200 import client
    with native.Service(["--port", "0", "--bind", "127.0.0.1", "--threads", "1", "--quiet"]) as svc:
        assert svc.port > 0
        for _ in range(3):
            assert client.get("127.0.0.1", svc.port, "/api/hola").body == b"Hola de unknown"
        assert svc.stats()["requests"] == 3
    with pytest.raises(native.NativeError, match="unknown argument"):
        native.Service(["--nope"]).start()
