"""Packaging and deployment artefacts and the cold-start measurement
(BASELINE.json config 3). The image and the manifest cannot be exercised on a box without a
container runtime or a cluster; what can be checked is that they say what the reference's say."""
from __future__ import annotations

import json
import os
import re
import socket
import subprocess
from pathlib import Path

import hola_b200
from hola_b200 import loadgen

REPO = Path(__file__).resolve().parent.parent


def free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_dockerfile_mirrors_the_reference_image():
    text = (REPO / "Dockerfile").read_text()
    assert re.search(r"^EXPOSE 8080$", text, re.M)  # Dockerfile:8
    assert re.search(r"^ENV ZIPKIN_SERVER_URL[= ]http://zipkin-query:9411$", text, re.M)  # Dockerfile:6
    assert "FROM scratch" in text and "hola-static" in text
    assert re.search(r'^ENTRYPOINT \["/app/hola"\]$', text, re.M)  # jar lived under /app too (Dockerfile:10)


def test_kubernetes_manifest_mirrors_the_fabric8_properties():
    doc = json.loads((REPO / "deploy" / "kubernetes.json").read_text())
    items = {i["kind"]: i for i in doc["items"]}
    svc, dep = items["Service"], items["Deployment"]
    assert svc["metadata"]["name"] == "hola"  # pom.xml:77
    port = svc["spec"]["ports"][0]
    assert (port["port"], port["targetPort"]) == (8080, 8080)  # pom.xml:78-79
    assert svc["spec"]["type"] == "LoadBalancer"  # pom.xml:80
    labels = dep["spec"]["template"]["metadata"]["labels"]
    assert labels["group"] == "msa" and labels["hystrix.enabled"] == "true"  # pom.xml:81-82
    container = dep["spec"]["template"]["spec"]["containers"][0]
    assert container["image"] == "redhatmsa/hola:0.0.1-SNAPSHOT"  # pom.xml:71-72
    env = {e["name"]: e.get("value") for e in container["env"]}
    assert env["AB_OFF"] == "true" and env["JAVA_OPTIONS"] == "-Xmx512m"  # pom.xml:88-89
    probe = container["readinessProbe"]
    assert probe["httpGet"] == {"path": "/api/health", "port": 8080}  # pom.xml:91-92
    assert probe["initialDelaySeconds"] == 5 and probe["timeoutSeconds"] == 30  # pom.xml:93-94
    for selector_key, value in svc["spec"]["selector"].items():
        assert labels[selector_key] == value  # the service actually selects the pods


def test_sample_config_files_are_loadable(hola_server):
    from hola_b200 import client, native
    entries = dict(native.properties_parse((REPO / "translation.properties").read_bytes()))
    assert entries == {b"hello": b"Hola de %s mediante el archivo de configuraciones\n"}
    hystrix = dict(native.properties_parse((REPO / "conf" / "config.properties").read_bytes()))
    assert hystrix[b"hystrix.command.default.execution.isolation.thread.timeoutInMilliseconds"] == b"600"
    assert hystrix[b"hystrix.command.default.circuitBreaker.requestVolumeThreshold"] == b"5"
    srv = hola_server(args=["--conf", str(REPO / "translation.properties"), "--hystrix-config", str(REPO / "conf" / "config.properties")])
    assert client.get("127.0.0.1", srv.port, "/api/hola").body == b"Hola de unknown mediante el archivo de configuraciones\n"


def test_static_executable_is_really_static():
    p = hola_b200.build("static")
    out = subprocess.run(["ldd", str(p.hola)], capture_output=True, text=True)
    assert "not a dynamic executable" in (out.stdout + out.stderr)
    # No libc resolver entry points: the service carries its own (src/net/dns.cc).
    syms = subprocess.run(["nm", str(p.hola)], capture_output=True, text=True).stdout
    assert " T getaddrinfo" not in syms and " T gethostbyname" not in syms


def test_cold_start_is_measured_in_milliseconds():
    port = free_port()
    env = {k: v for k, v in os.environ.items() if k not in ("HOSTNAME", "ZIPKIN_SERVER_URL")}
    rep = loadgen.cold_start([str(hola_b200.paths().hola), "--port", str(port), "--bind", "127.0.0.1", "--threads", "2"],
                             port=port, runs=5, env=env)
    assert len(rep["runs"]) == 5 and rep["unit"] == "us"
    assert 100 < rep["median"] < 500_000  # the JVM service is given 5 s before its first probe (pom.xml:93)
