"""The built-in stub resolver (src/net/dns.cc) against a fake DNS server: the downstream service is
addressed by name ("aloha", HolaResource.java:95,103) and in Kubernetes that name only resolves
through the pod's search list."""
from __future__ import annotations

import json
import socket
import struct
import threading
import time

import pytest

from hola_b200 import client

FALLBACK = "Aloha response (fallback)"


from hola_b200.fakes import FakeDns  # noqa: E402


@pytest.fixture
def fake_dns():
    servers = []

    def make(records):
        d = FakeDns(records)
        servers.append(d)
        return d

    yield make
    for d in servers:
        d.stop()


def dns_env(tmp_path, dns, search="msa.svc.cluster.local svc.cluster.local", options="ndots:5 timeout:1 attempts:1"):
    conf = tmp_path / "resolv.conf"
    conf.write_text("nameserver 127.0.0.1\nsearch %s\noptions %s\n" % (search, options))
    hosts = tmp_path / "hosts"
    hosts.write_text("127.0.0.1 localhost\n")
    return {"HOLA_RESOLV_CONF": str(conf), "HOLA_HOSTS_FILE": str(hosts), "HOLA_DNS_PORT": str(dns.port)}


def chain(srv):
    r = client.get("127.0.0.1", srv.port, "/api/hola-chaining")
    assert r.status == 200
    return json.loads(r.body)


def test_service_name_resolves_through_the_search_list(hola_server, fake_aloha, fake_dns, tmp_path):
    dns = fake_dns({"aloha.msa.svc.cluster.local": "127.0.0.1"})
    srv = hola_server(args=["--aloha-url", "http://aloha:%d/" % fake_aloha.port], env=dns_env(tmp_path, dns))
    assert chain(srv) == ["Hola de unknown"] + fake_aloha.payload
    assert dns.queries[0] == "aloha.msa.svc.cluster.local"
    # Positive answers are cached (30 s, like the JVM): no second query.
    n = len(dns.queries)
    assert chain(srv)[1:] == fake_aloha.payload
    assert len(dns.queries) == n
    assert fake_aloha.snapshot()[0]["headers"]["host"] == "aloha:%d" % fake_aloha.port


def test_second_search_domain_and_bare_name(hola_server, fake_aloha, fake_dns, tmp_path):
    dns = fake_dns({"aloha.svc.cluster.local": "127.0.0.1"})
    srv = hola_server(args=["--aloha-url", "http://aloha:%d/" % fake_aloha.port], env=dns_env(tmp_path, dns))
    assert chain(srv)[1:] == fake_aloha.payload
    assert dns.queries[:2] == ["aloha.msa.svc.cluster.local", "aloha.svc.cluster.local"]
    dns2 = fake_dns({"aloha": "127.0.0.1"})
    srv2 = hola_server(args=["--aloha-url", "http://aloha:%d/" % fake_aloha.port], env=dns_env(tmp_path, dns2))
    assert chain(srv2)[1:] == fake_aloha.payload
    assert dns2.queries == ["aloha.msa.svc.cluster.local", "aloha.svc.cluster.local", "aloha"]


def test_unknown_name_falls_back_and_is_negatively_cached(hola_server, fake_dns, tmp_path):
    dns = fake_dns({})
    srv = hola_server(args=["-Dhystrix.command.default.circuitBreaker.enabled=false"], env=dns_env(tmp_path, dns))
    assert chain(srv) == ["Hola de unknown", FALLBACK]
    n = len(dns.queries)
    assert n == 3
    assert chain(srv) == ["Hola de unknown", FALLBACK]
    assert len(dns.queries) == n  # 10 s negative cache, like the JVM


def test_slow_dns_is_cut_off_by_the_command_timeout(hola_server, fake_aloha, fake_dns, tmp_path):
    dns = fake_dns({"aloha.msa.svc.cluster.local": "127.0.0.1"})
    dns.silent = True
    env = dns_env(tmp_path, dns, options="ndots:5 timeout:5 attempts:2")
    srv = hola_server(args=["--aloha-url", "http://aloha:%d/" % fake_aloha.port], env=env)
    t0 = time.time()
    assert chain(srv) == ["Hola de unknown", FALLBACK]
    assert 0.55 < time.time() - t0 < 1.2  # the 600 ms Hystrix timeout, not the resolver's 10 s
    # Meanwhile the service keeps answering other requests.
    assert client.get("127.0.0.1", srv.port, "/api/hola").status == 200
    t0 = time.time()
    assert srv.stop() == 0
    assert time.time() - t0 < 3  # a lookup stuck on a dead name server does not hold up shutdown


def test_answer_to_a_different_question_is_ignored(hola_server, fake_aloha, fake_dns, tmp_path):
    """A datagram with the right id but another question section (a spoofing attempt, or a late
    answer to an earlier query) is not the answer: the resolver keeps waiting for the real one."""
    dns = fake_dns({"aloha.msa.svc.cluster.local": "127.0.0.1"})
    dns.stray_first = ("evil.example", "127.0.0.9")  # nothing listens there: taking it would mean the fallback
    srv = hola_server(args=["--aloha-url", "http://aloha:%d/" % fake_aloha.port, "--feign-max-attempts", "1"],
                      env=dns_env(tmp_path, dns))
    assert chain(srv) == ["Hola de unknown"] + fake_aloha.payload


def test_hosts_file_beats_dns(hola_server, fake_aloha, fake_dns, tmp_path):
    dns = fake_dns({})
    env = dns_env(tmp_path, dns)
    (tmp_path / "hosts").write_text("127.0.0.1 localhost\n127.0.0.1 aloha\n")
    srv = hola_server(args=["--aloha-url", "http://aloha:%d/" % fake_aloha.port], env=env)
    assert chain(srv)[1:] == fake_aloha.payload
    assert dns.queries == []


def test_zipkin_host_is_resolved_the_same_way(hola_server, fake_zipkin, fake_dns, tmp_path):
    dns = fake_dns({"zipkin-query.msa.svc.cluster.local": "127.0.0.1"})
    env = dns_env(tmp_path, dns)
    env["ZIPKIN_SERVER_URL"] = "http://zipkin-query:%d" % fake_zipkin.port  # Dockerfile:6 uses this host name
    srv = hola_server(args=["--zipkin-flush-ms", "100"], env=env)
    client.get("127.0.0.1", srv.port, "/api/hola")
    assert len(fake_zipkin.wait_for_spans(1)) == 1
    assert "zipkin-query.msa.svc.cluster.local" in dns.queries


def test_resolve_flag_pins_a_name(hola_server, fake_aloha):
    srv = hola_server(args=["--aloha-url", "http://aloha:%d/" % fake_aloha.port, "--resolve", "aloha=127.0.0.1"])
    assert chain(srv)[1:] == fake_aloha.payload


def test_truncated_udp_answer_is_retried_over_tcp(hola_server, fake_aloha, fake_dns, tmp_path):
    dns = fake_dns({"aloha.msa.svc.cluster.local": "127.0.0.1"})
    dns.truncate_udp = True
    srv = hola_server(args=["--aloha-url", "http://aloha:%d/" % fake_aloha.port], env=dns_env(tmp_path, dns))
    assert chain(srv)[1:] == fake_aloha.payload
    assert dns.tcp_queries == ["aloha.msa.svc.cluster.local"]


def test_second_name_server_is_tried_when_the_first_is_dead(hola_server, fake_aloha, fake_dns, tmp_path):
    dns = fake_dns({"aloha.msa.svc.cluster.local": "127.0.0.1"})
    env = dns_env(tmp_path, dns, options="ndots:5 timeout:1 attempts:2")
    # 127.0.0.2 answers nothing on that port (ICMP unreachable on loopback), the second entry is the fake.
    (tmp_path / "resolv.conf").write_text("nameserver 127.0.0.2\nnameserver 127.0.0.1\nsearch msa.svc.cluster.local\n"
                                          "options ndots:5 timeout:1 attempts:2\n")
    srv = hola_server(args=["--aloha-url", "http://aloha:%d/" % fake_aloha.port,
                            "-Dhystrix.command.default.execution.isolation.thread.timeoutInMilliseconds=3000"], env=env)
    assert chain(srv)[1:] == fake_aloha.payload


class _V6Aloha:
    """A downstream that exists only on IPv6 (an IPv6-only cluster)."""

    def __init__(self, payload):
        from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
        seen = self.seen = []

        class Handler(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"

            def log_message(self, *a):
                pass

            def do_GET(self):
                seen.append(self.headers.get("Host"))
                body = json.dumps(payload).encode()
                self.send_response(200)
                self.send_header("Content-Type", "application/json")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

        class Server(ThreadingHTTPServer):
            address_family = socket.AF_INET6
            daemon_threads = True

        self.httpd = Server(("::1", 0), Handler)
        self.port = self.httpd.server_address[1]
        threading.Thread(target=self.httpd.serve_forever, daemon=True).start()

    def stop(self):
        self.httpd.shutdown()
        self.httpd.server_close()


@pytest.fixture
def v6_aloha():
    try:
        srv = _V6Aloha(["Aloha over IPv6"])
    except OSError:
        pytest.skip("no IPv6 loopback on this machine")
    yield srv
    srv.stop()


def test_ipv6_only_downstream_via_aaaa(hola_server, v6_aloha, fake_dns, tmp_path):
    """The name has no A record, only AAAA: the second question finds it."""
    dns = fake_dns({"aloha.msa.svc.cluster.local": "::1"})
    srv = hola_server(args=["--aloha-url", "http://aloha:%d/" % v6_aloha.port], env=dns_env(tmp_path, dns))
    assert chain(srv) == ["Hola de unknown", "Aloha over IPv6"]
    assert dns.queries[:2] == ["aloha.msa.svc.cluster.local", "aloha.msa.svc.cluster.local"]  # A, then AAAA


def test_ipv6_literal_and_pinned_downstream(hola_server, v6_aloha):
    srv = hola_server(args=["--aloha-url", "http://[::1]:%d/" % v6_aloha.port])
    assert chain(srv) == ["Hola de unknown", "Aloha over IPv6"]
    assert v6_aloha.seen[-1] == "[::1]:%d" % v6_aloha.port  # literal goes into the Host header in brackets
    srv2 = hola_server(args=["--aloha-url", "http://aloha:%d/" % v6_aloha.port, "--resolve", "aloha=::1"])
    assert chain(srv2) == ["Hola de unknown", "Aloha over IPv6"]
