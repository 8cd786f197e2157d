"""The observable contract of every endpoint, byte for byte where the reference's
own code determines the bytes."""
from __future__ import annotations

import json
import os
import time

import pytest

from hola_b200 import client

DATA = os.path.join(os.path.dirname(__file__), "data")
CONF = os.path.join(DATA, "translation.properties")
REFERENCE_CONF = "/root/reference/translation.properties"


HOLA = b"Hola de unknown"


def get(srv, target, **kw):
    return client.get("127.0.0.1", srv.port, target, **kw)


# ---- GET /api/hola: the three cases of SURVEY.md 2.2 ------------------------------------------

def test_hola_default_no_hostname(hola_server):
    """No -Dconf, HOSTNAME not exported: `Hola de unknown`, 15 bytes, no newline (HolaResource.java:57-65)."""
    srv = hola_server()
    r = get(srv, "/api/hola")
    assert r.status == 200 and r.reason == "OK" and r.version == "HTTP/1.1"
    assert r.body == b"Hola de unknown" and len(r.body) == 15
    assert r.header("Content-Type") == "text/plain"
    assert r.header("Content-Length") == "15"
    assert r.header("Access-Control-Allow-Origin") == "*"  # CorsFilter.java:31
    assert r.header("Date") is not None


def test_hola_hostname_from_env(hola_server):
    srv = hola_server(env={"HOSTNAME": "hola-1-x7k2p"})
    assert get(srv, "/api/hola").body == b"Hola de hola-1-x7k2p"


def test_hola_empty_hostname_is_kept(hola_server):
    """getOrDefault only substitutes when the variable is absent, not when it is empty."""
    srv = hola_server(env={"HOSTNAME": ""})
    assert get(srv, "/api/hola").body == b"Hola de "


def test_hola_with_conf_file(hola_server):
    """-Dconf=translation.properties: the value's trailing backslash-n becomes one LF (BASELINE.json config 2)."""
    srv = hola_server(args=["-Dconf=" + CONF], env={"HOSTNAME": "pod-7"})
    r = get(srv, "/api/hola")
    assert r.body == b"Hola de pod-7 mediante el archivo de configuraciones\n"
    assert r.body.count(b"\n") == 1


@pytest.mark.skipif(not os.path.exists(REFERENCE_CONF), reason="reference checkout not mounted")
def test_hola_with_the_reference_file_itself(hola_server):
    srv = hola_server(args=["--conf", REFERENCE_CONF])
    assert get(srv, "/api/hola").body == b"Hola de unknown mediante el archivo de configuraciones\n"


def test_hola_conf_file_missing_or_broken_falls_back(hola_server, tmp_path):
    """Any failure reading the file means "no entries" (Configuration.java:34-36)."""
    srv = hola_server(args=["-Dconf=" + str(tmp_path / "absent.properties")])
    assert get(srv, "/api/hola").body == b"Hola de unknown"
    broken = tmp_path / "broken.properties"
    broken.write_text("hello=caf\\u00e\n")
    srv2 = hola_server(args=["-Dconf=" + str(broken)])
    assert get(srv2, "/api/hola").body == b"Hola de unknown"


def test_hola_hot_reload_within_five_seconds(hola_server, tmp_path):
    """Edits to the file show up within the 5 s cache period, without a restart (HolaResource.java:63)."""
    conf = tmp_path / "t.properties"
    conf.write_text("hello=Uno de %s\n")
    srv = hola_server(args=["-Dconf=" + str(conf)])
    assert get(srv, "/api/hola").body == b"Uno de unknown"
    conf.write_text("hello=Dos de %s\n")
    assert get(srv, "/api/hola").body == b"Uno de unknown"  # still inside the cache period
    deadline = time.time() + 6.5
    body = b""
    while time.time() < deadline:
        body = get(srv, "/api/hola").body
        if body == b"Dos de unknown":
            break
        time.sleep(0.1)
    assert body == b"Dos de unknown"
    # logChanges(true), HolaResource.java:61. The line travels through a pipe and a reader thread:
    # give it a moment.
    deadline = time.time() + 3
    while "New value Dos de %s for key hello." not in srv.stderr and time.time() < deadline:
        time.sleep(0.05)
    assert "New value Dos de %s for key hello." in srv.stderr


def test_hola_source_precedence(hola_server, tmp_path):
    """System property (400) beats environment (300) beats the conf file (0)."""
    conf = tmp_path / "t.properties"
    conf.write_text("hello=file %s\n")
    assert get(hola_server(args=["-Dconf=" + str(conf)]), "/api/hola").body == b"file unknown"
    assert get(hola_server(args=["-Dconf=" + str(conf)], env={"hello": "env %s"}), "/api/hola").body == b"env unknown"
    srv = hola_server(args=["-Dconf=" + str(conf), "-Dhello=prop %s"], env={"hello": "env %s"})
    assert get(srv, "/api/hola").body == b"prop unknown"


def test_hola_pattern_the_jvm_would_reject_is_a_500(hola_server):
    srv = hola_server(args=["-Dhello=Hola %s y %s"])  # MissingFormatArgumentException on the JVM
    r = get(srv, "/api/hola")
    assert r.status == 500
    srv2 = hola_server(args=["-Dhello=100%% de %s%n"])
    assert get(srv2, "/api/hola").body == b"100% de unknown\n"


def test_hola_log_config_reads_prints_path(hola_server, tmp_path):
    """The reference prints the conf path on every read (Configuration.java:28); opt-in here."""
    conf = tmp_path / "t.properties"
    conf.write_text("hello=x %s\n")
    srv = hola_server(args=["-Dconf=" + str(conf), "--log-config-reads"])
    get(srv, "/api/hola")
    srv.stop()
    assert str(conf) in srv.stdout
    srv2 = hola_server(args=["--log-config-reads"])
    get(srv2, "/api/hola")
    srv2.stop()
    assert "null" in srv2.stdout  # System.out.println((String) null)


# ---- the other endpoints -----------------------------------------------------------------------

def test_health(hola_server):
    srv = hola_server()
    r = get(srv, "/api/health")
    assert r.status == 200 and r.body == b"I'm ok" and len(r.body) == 6  # HolaResource.java:85
    assert r.header("Content-Type") == "text/plain"
    assert r.header("Access-Control-Allow-Origin") == "*"


def test_head_and_options_and_405(hola_server):
    srv = hola_server()
    r = get(srv, "/api/hola", method="HEAD")
    assert r.status == 200 and r.body == b"" and r.header("Content-Length") == "15"
    r = get(srv, "/api/hola", method="OPTIONS")
    assert r.status == 200 and "GET" in r.header("Allow")
    for method in ("POST", "PUT", "DELETE", "PATCH"):
        r = get(srv, "/api/hola", method=method)
        assert r.status == 405 and "GET" in r.header("Allow"), method
        assert r.header("Access-Control-Allow-Origin") == "*"


def test_not_found_and_not_acceptable(hola_server):
    srv = hola_server()
    r = get(srv, "/api/nothing")
    assert r.status == 404 and r.header("Access-Control-Allow-Origin") == "*"
    r = get(srv, "/elsewhere")
    assert r.status == 404 and r.header("Access-Control-Allow-Origin") is None  # CORS is a JAX-RS filter: /api only
    r = get(srv, "/api/hola", headers={"Accept": "application/json"})
    assert r.status == 406
    r = get(srv, "/api/hola", headers={"Accept": "text/html, */*;q=0.5"})
    assert r.status == 200
    r = get(srv, "/api/hola-chaining", headers={"Accept": "text/plain"})
    assert r.status == 406


def test_error_page_does_not_reflect_markup(hola_server):
    srv = hola_server()
    r = get(srv, "/api/<script>alert(1)</script>", headers={"Host": "evil\"><img src=x>"})
    assert r.status == 404
    assert b"<script>" not in r.body and b"<img" not in r.body
    assert b"&lt;script&gt;" in r.body


def test_query_string_and_trailing_slash(hola_server):
    srv = hola_server()
    assert get(srv, "/api/hola?x=1&y=2").body == b"Hola de unknown"
    assert get(srv, "/api/hola/").status == 200
    assert get(srv, "/api/holaa").status == 404
    assert get(srv, "/api/hol").status == 404
    for bare in ("/api", "/api/"):
        r = get(srv, bare)
        assert r.status == 404 and r.header("Access-Control-Allow-Origin") == "*"  # still inside the JAX-RS mapping
    assert get(srv, "/apix").header("Access-Control-Allow-Origin") is None


def test_landing_page_is_the_reference_page(hola_server):
    srv = hola_server()
    r = get(srv, "/")
    assert r.status == 200 and r.header("Content-Type").startswith("text/html")
    assert len(r.body) == 485
    assert not r.body.endswith(b"\n")
    assert b"webjars/swagger-ui/2.1.4/?url=' + loc + 'api/swagger.json#!/default'" in r.body
    ref = "/root/reference/src/main/webapp/index.html"
    if os.path.exists(ref):
        assert r.body == open(ref, "rb").read()
    assert r.header("Access-Control-Allow-Origin") is None
    assert get(srv, "/index.html").body == r.body


def test_swagger_document(hola_server):
    srv = hola_server()
    r = get(srv, "/api/swagger.json")
    assert r.status == 200 and r.header("Content-Type") == "application/json"
    assert r.header("Access-Control-Allow-Origin") == "*"
    doc = json.loads(r.body)
    assert doc["swagger"] == "2.0"
    # SwaggerListener.java:31-40
    assert doc["info"]["version"] == "1.0.0"
    assert doc["info"]["title"] == "Hola microservices REST API"
    assert doc["info"]["description"] == "Operations that can be invoked in the hola microservices"
    assert doc["info"]["license"] == {"name": "Apache 2.0", "url": "http://www.apache.org/licenses/LICENSE-2.0.html"}
    assert doc["info"]["contact"]["name"] == "developer@redhat.com"
    assert doc["basePath"] == "/api" and doc["schemes"] == ["http"]
    # @ApiOperation values, HolaResource.java:55,72,83
    assert doc["paths"]["/hola"]["get"]["summary"] == "Returns the greeting in Spanish"
    assert doc["paths"]["/hola-chaining"]["get"]["summary"] == "Returns the greeting plus the next service in the chain"
    assert doc["paths"]["/health"]["get"]["summary"] == "Used to verify the health of the service"
    assert doc["paths"]["/hola"]["get"]["produces"] == ["text/plain"]
    assert doc["paths"]["/hola-chaining"]["get"]["produces"] == ["application/json"]
    assert doc["paths"]["/hola-chaining"]["get"]["responses"]["200"]["schema"] == {"type": "array", "items": {"type": "string"}}
    assert b'"swagger" : "2.0"' in r.body  # setPrettyPrint(true)
    y = get(srv, "/api/swagger.yaml")
    assert y.status == 200 and b'swagger: "2.0"' in y.body and b"/hola-chaining:" in y.body


def test_request_path_is_canonicalised_before_routing(hola_server, fake_zipkin):
    """Undertow decodes and normalises the request path, and RESTEasy matches it without matrix
    parameters (JaxRsActivator.java:30, HolaResource.java:46,52-53,69-70,80-81): these spellings
    all name the same three resources. The path as sent survives where the reference would show
    it: in RESTEasy's 404 text and in the span's http.url tag."""
    srv = hola_server(env={"ZIPKIN_SERVER_URL": fake_zipkin.url}, args=["--zipkin-flush-ms", "50"])
    for target in ("/api/%68ola", "/api/hola;a=1", "/api/../api/hola", "/api//hola", "/api/./hola", "/api;v=2/hola", "/%61pi/hol%61",
                   "/api/x/../hola", "//api/hola", "/api/%2e%2e/api/hola"):
        r = get(srv, target)
        assert (r.status, r.body) == (200, HOLA), target
        assert r.header("Access-Control-Allow-Origin") == "*"
    assert get(srv, "/api/../api/health").body == b"I'm ok"
    assert get(srv, "/api/%68ealth;probe=1?x=1").body == b"I'm ok"
    assert get(srv, "/api/hola-chaining/../swagger.json").header("Content-Type") == "application/json"
    assert get(srv, "/./index.html").status == 200 and get(srv, "/x/..").status == 200  # the landing page
    # Not the same resource: an encoded slash is no separator, an encoded semicolon no parameter.
    assert get(srv, "/api%2Fhola").status == 404
    assert get(srv, "/api/hola%3Ba=1").status == 404
    assert get(srv, "/api/hola/extra/..;x").status == 200  # "..;x" is ".." once its parameters are cut
    # Malformed escapes and NUL bytes are refused outright.
    for target in ("/api/hola%", "/api/hola%4", "/api/hola%zz", "/api/hola%00"):
        r = get(srv, target)
        assert r.status == 400 and r.header("Connection") == "close", target
    # The 404 text quotes the URL as it was sent.
    r = get(srv, "/api/no%74here;p=1")
    assert r.status == 404 and b"/api/no%74here;p=1" in r.body
    # Tracing stays mapped on the canonical URL (ZipKinFilter.java:43) and reports the URL as sent.
    urls = sorted(s["binaryAnnotations"][0]["value"] for s in fake_zipkin.wait_for_spans(12))
    assert "http://localhost/api/%68ola" in urls and "http://localhost/api/hola;a=1" in urls
    assert not any("health" in u or "swagger" in u for u in urls)


def test_webjars_directory(hola_server, tmp_path):
    root = tmp_path / "webjars"
    ui = root / "swagger-ui" / "2.1.4"
    ui.mkdir(parents=True)
    (ui / "index.html").write_text("<html>swagger ui</html>")
    (ui / "swagger-ui.js").write_text("var x = 1;")
    (tmp_path / "secret.txt").write_text("top secret")
    srv = hola_server(args=["--webjars-dir", str(root)])
    r = get(srv, "/webjars/swagger-ui/2.1.4/?url=http://x/api/swagger.json")
    assert r.status == 200 and r.body == b"<html>swagger ui</html>" and r.header("Content-Type") == "text/html"
    r = get(srv, "/webjars/swagger-ui/2.1.4/swagger-ui.js")
    assert r.status == 200 and r.header("Content-Type") == "application/javascript"
    assert get(srv, "/webjars/swagger-ui/2.1.4/missing.css").status == 404
    assert get(srv, "/webjars/../secret.txt").status == 404
    assert get(srv, "/webjars/%2e%2e/secret.txt").status == 404
    assert get(srv, "/webjars/swagger-ui/2.1.4/..%2f..%2f..%2fsecret.txt").status == 404
    # A symbolic link inside the directory must not lead out of it; one that stays inside is fine.
    (ui / "link.txt").symlink_to(tmp_path / "secret.txt")
    (ui / "alias.js").symlink_to(ui / "swagger-ui.js")
    (root / "up").symlink_to(tmp_path)
    assert get(srv, "/webjars/swagger-ui/2.1.4/link.txt").status == 404
    assert get(srv, "/webjars/up/secret.txt").status == 404
    assert get(srv, "/webjars/swagger-ui/2.1.4/alias.js").body == b"var x = 1;"
    # Served from memory, but a changed file is noticed (size or mtime).
    assert get(srv, "/webjars/swagger-ui/2.1.4/swagger-ui.js").body == b"var x = 1;"
    (ui / "swagger-ui.js").write_text("var x = 22;")
    assert get(srv, "/webjars/swagger-ui/2.1.4/swagger-ui.js").body == b"var x = 22;"
    r = client.get("127.0.0.1", srv.port, "/webjars/swagger-ui/2.1.4/swagger-ui.js", method="HEAD")
    assert r.status == 200 and r.header("Content-Length") == "11"
    # Without --webjars-dir the Swagger-UI assets are absent (they are not on the build box either),
    # but the page the landing page frames is replaced by a built-in API browser, so "/" works.
    bare = hola_server()
    for target in ("/webjars/swagger-ui/2.1.4/?url=http://localhost:8080/api/swagger.json", "/webjars/swagger-ui/2.1.4/index.html"):
        r = get(bare, target)
        assert r.status == 200 and r.header("Content-Type") == "text/html"
        assert b"Try it out" in r.body and b"fetch(specUrl)" in r.body
        assert b"src=\"http" not in r.body and b"href=\"http" not in r.body  # self-contained: no external assets
    assert get(bare, "/webjars/swagger-ui/2.1.4/swagger-ui.js").status == 404
    assert get(bare, "/webjars/other/1.0/x.js").status == 404


def test_cli_version_help_and_bad_args(native_build):
    import subprocess
    exe = str(native_build.hola)
    out = subprocess.run([exe, "--version"], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.startswith("hola 0.0.1-SNAPSHOT")
    out = subprocess.run([exe, "--help"], capture_output=True, text=True)
    assert out.returncode == 0 and "-Dconf=FILE" in out.stdout
    out = subprocess.run([exe, "--bogus"], capture_output=True, text=True)
    assert out.returncode == 2 and "unknown argument" in out.stderr
    out = subprocess.run([exe, "--port", "99999"], capture_output=True, text=True)
    assert out.returncode == 2
    # --print-greeting: the /api/hola body for a configuration, without starting a server.
    env = {k: v for k, v in os.environ.items() if k not in ("HOSTNAME", "hello")}
    out = subprocess.run([exe, "--print-greeting", "-Dconf=" + CONF], capture_output=True, env=dict(env, HOSTNAME="pod-3"))
    assert out.returncode == 0 and out.stdout == b"Hola de pod-3 mediante el archivo de configuraciones\n"
    out = subprocess.run([exe, "--print-greeting"], capture_output=True, env=env)
    assert out.returncode == 0 and out.stdout == b"Hola de unknown"
    out = subprocess.run([exe, "--print-greeting", "-Dhello=%d items"], capture_output=True, env=env)
    assert out.returncode == 1 and b"IllegalFormatConversionException" in out.stderr


def test_port_in_use_is_reported(hola_server, native_build):
    import subprocess
    srv = hola_server(args=["--accept-mode", "rr"])  # no SO_REUSEPORT, so a second bind must fail
    out = subprocess.run([str(native_build.hola), "--port", str(srv.port), "--bind", "127.0.0.1", "--accept-mode", "rr",
                          "--threads", "1"], capture_output=True, text=True, timeout=10)
    assert out.returncode == 1 and "bind" in out.stderr


def test_swarm_style_port_definition(hola_server):
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    srv = hola_server(args=["-Dswarm.http.port=%d" % port])
    assert srv.port == port
    assert get(srv, "/api/health").status == 200
