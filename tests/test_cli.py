"""The command line of the executable: every option's error path, the informational modes, the
shutdown profile and the counter dump. The reference is started as
``java -jar hola-swarm.jar [-Dconf=...] [-Dhello=...]`` (/root/reference/Readme.md:13-16,
Configuration.java:27); the options beyond the -D definitions are this service's own."""
from __future__ import annotations

import json
import os
import signal
import subprocess
import time

import pytest

import hola_b200
from hola_b200 import client

CONF = str(hola_b200.REPO_ROOT / "tests" / "data" / "translation.properties")


@pytest.fixture
def exe(request):
    """The executable of the build variant under test (pytest --variant ...)."""
    variant = request.config.getoption("--variant")
    if variant != "release":
        hola_b200.build(variant)
    return str(hola_b200.paths(variant).hola)


def clean_env(**extra):
    env = {k: v for k, v in os.environ.items() if k not in ("HOSTNAME", "hello", "ZIPKIN_SERVER_URL")}
    env.update(extra)
    return env


def run(exe, *args, **kw):
    return subprocess.run([exe, *args], capture_output=True, text=True, timeout=20, env=kw.pop("env", clean_env()), **kw)


USAGE_ERRORS = [
    (["--conf"], "--conf needs a value"),
    (["--port"], "--port needs a value"),
    (["--port", "99999"], "bad port"),
    (["--port", "80a"], "bad port"),
    (["-Dswarm.http.port=-1"], "bad port"),
    (["--bind"], "--bind needs a value"),
    (["--threads"], "--threads needs a number"),
    (["--threads", "many"], "--threads needs a number"),
    (["--threads", "99999"], "--threads needs a number"),
    (["--cpus"], "bad --cpus list"),
    (["--cpus", "0,x"], "bad --cpus list"),
    (["--cpus", "5-2"], "bad --cpus list"),
    (["--cpus", ","], "bad --cpus list"),
    (["--accept-mode", "random"], "--accept-mode is balanced, reuseport or rr"),
    (["--io-backend", "iocp"], "--io-backend is auto, epoll or uring"),
    (["--aloha-url", "https://aloha/"], "--aloha-url must look like"),
    (["--aloha-url"], "--aloha-url must look like"),
    (["--feign-max-attempts", "0"], "--feign-max-attempts needs a number >= 1"),
    (["--resolve"], "--resolve needs NAME=ADDRESS"),
    (["--resolve", "aloha"], "--resolve needs NAME=ADDRESS"),
    (["--resolve", "aloha=not-an-address"], "--resolve needs NAME=ADDRESS"),
    (["--resolve", "=127.0.0.1"], "--resolve needs NAME=ADDRESS"),
    (["--zipkin-url"], "--zipkin-url needs a value"),
    (["--zipkin-flush-ms", "0"], "--zipkin-flush-ms needs a positive number"),
    (["--zipkin-ring", "1"], "--zipkin-ring needs a number of spans >= 2"),
    (["--zipkin-max-batch", "0"], "--zipkin-max-batch needs a positive number"),
    (["--hystrix-config"], "--hystrix-config needs a value"),
    (["--hystrix-config", "/no/such/file.properties"], "--hystrix-config:"),
    (["--deltaspike-properties"], "--deltaspike-properties needs a value"),
    (["--webjars-dir"], "--webjars-dir needs a value"),
    (["--idle-timeout", "soon"], "--idle-timeout needs milliseconds"),
    (["--drain-linger-ms", "-5"], "--drain-linger-ms needs milliseconds"),
    (["--spin-us", "lots"], "--spin-us needs microseconds"),
    (["--max-header-bytes", "100"], "--max-header-bytes needs a number >= 256"),
    (["--bogus"], "unknown argument: --bogus"),
]


@pytest.mark.parametrize("args,message", USAGE_ERRORS, ids=[" ".join(a) for a, _ in USAGE_ERRORS])
def test_usage_errors_exit_2_with_a_reason(exe, args, message):
    out = run(exe, *args)
    assert out.returncode == 2, out.stderr
    assert message in out.stderr and out.stdout == ""


def test_version_help_and_print_greeting(exe, tmp_path):
    out = run(exe, "--version")
    assert out.returncode == 0 and out.stdout == "hola 0.0.1-SNAPSHOT\n"
    for flag in ("--help", "-h"):
        out = run(exe, flag)
        assert out.returncode == 0 and out.stdout.startswith("usage: hola") and "--io-backend" in out.stdout and out.stderr == ""
    out = run(exe, "--print-greeting", "--conf", CONF, env=clean_env(HOSTNAME="pod-3"))
    assert (out.returncode, out.stdout) == (0, "Hola de pod-3 mediante el archivo de configuraciones\n")
    out = run(exe, "--print-greeting", "-Dhello=%d items")
    assert out.returncode == 1 and "IllegalFormatConversionException" in out.stderr and out.stdout == ""
    # -Dhello without a value defines the empty string, and DeltaSpike's resolver falls back to the
    # default for an empty value (HolaResource.java:58-64 gives it "Hola de %s" as the default).
    out = run(exe, "--print-greeting", "-Dhello")
    assert (out.returncode, out.stdout) == (0, "Hola de unknown")
    # An extra properties file at ordinal 100 (META-INF/apache-deltaspike.properties): beats the conf file, loses to -D.
    extra = tmp_path / "apache-deltaspike.properties"
    extra.write_text("hello=Bonjour de %s\n")
    out = run(exe, "--print-greeting", "--deltaspike-properties", str(extra), "--conf", CONF)
    assert out.stdout == "Bonjour de unknown"
    out = run(exe, "--print-greeting", "--deltaspike-properties", str(extra), "-Dhello=Ciao %s")
    assert out.stdout == "Ciao unknown"


def test_a_port_that_cannot_be_bound_is_exit_1(exe, hola_server):
    srv = hola_server(args=["--accept-mode", "rr"])  # no SO_REUSEPORT, so a second bind must fail
    out = run(exe, "--port", str(srv.port), "--bind", "127.0.0.1", "--accept-mode", "rr", "--threads", "1")
    assert out.returncode == 1 and "bind 127.0.0.1:%d" % srv.port in out.stderr
    out = run(exe, "--bind", "not-an-address", "--port", "0")
    assert out.returncode == 1 and "bad bind address" in out.stderr


def test_every_option_is_accepted_together_and_verbose_prints_a_profile(exe, tmp_path, fake_aloha):
    hystrix = tmp_path / "config.properties"
    hystrix.write_text("hystrix.command.default.execution.isolation.thread.timeoutInMilliseconds=300\n")
    webjars = tmp_path / "webjars"
    webjars.mkdir()
    cpus = sorted(os.sched_getaffinity(0))
    cpu_list = "%d-%d" % (cpus[0], cpus[0]) if len(cpus) == 1 else "%d,%d-%d" % (cpus[0], cpus[1], cpus[1])
    proc = subprocess.Popen(
        [exe, "--port", "0", "--bind", "127.0.0.1", "--threads", "2", "--cpus", cpu_list, "--accept-mode", "reuseport", "--io-backend", "epoll",
         "--aloha-url", fake_aloha.url, "--feign-max-attempts", "2", "--resolve", "aloha=127.0.0.1", "--zipkin-url", "http://127.0.0.1:9",
         "--zipkin-flush-ms", "100", "--zipkin-ring", "64", "--zipkin-max-batch", "10", "--hystrix-config", str(hystrix),
         "--webjars-dir", str(webjars) + "//", "--idle-timeout", "60000", "--drain-linger-ms", "50", "--spin-us", "20",
         "--max-header-bytes", "4096", "--log-config-reads", "--log-feign", "--reference-logging", "--serve-aloha", "--verbose",
         "-Dhystrix.threadpool.default.coreSize=3", "-Dconf=" + CONF],
        stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=clean_env())
    try:
        line = ""
        while "listening on" not in line:
            line = proc.stderr.readline()
            assert line, "exited early"
        port = int(line.split("listening on 127.0.0.1:")[1].split()[0])
        assert "with 2 thread(s)" in line and "epoll" in line and "tracing on" in line
        assert client.get("127.0.0.1", port, "/api/hola").body == b"Hola de unknown mediante el archivo de configuraciones\n"
        assert json.loads(client.get("127.0.0.1", port, "/api/hola-chaining").body)[1:] == fake_aloha.payload
        assert client.get("127.0.0.1", port, "/api/aloha-chaining").body == b'["Aloha mai unknown"]'
        proc.send_signal(signal.SIGUSR1)
        time.sleep(0.2)
    finally:
        proc.send_signal(signal.SIGTERM)
        out, err = proc.communicate(timeout=15)
    assert proc.returncode == 0
    assert CONF in out  # --log-config-reads: the path, on stdout, as Configuration.java:28 prints it
    stats = json.loads([ln for ln in err.splitlines() if ln.startswith("hola-stats ")][0][len("hola-stats "):])
    assert stats["requests"] == 3 and stats["accept_mode"] == "reuseport" and stats["io_backend"] == "epoll" and len(stats["workers"]) == 2
    assert "[AlohaService#aloha] ---> GET" in err and "[AlohaService#aloha] <--- HTTP/1.1 200" in err  # --log-feign
    assert "signal 15, shutting down" in err
    profile = [ln for ln in err.splitlines() if "stats: requests=3" in ln]
    assert profile and "us_per_request(user/sys)=" in profile[0] and "spans(reported/dropped)=" in profile[0]
    workers = [ln for ln in err.splitlines() if "workers (requests/connections/cpu_s):" in ln]
    assert workers and " 0=" in workers[0] and " 1=" in workers[0]


def test_quiet_prints_nothing_on_a_clean_run(exe):
    proc = subprocess.Popen([exe, "--port", "0", "--bind", "127.0.0.1", "--threads", "1", "--quiet"], stdout=subprocess.PIPE,
                            stderr=subprocess.PIPE, text=True, env=clean_env())
    time.sleep(0.3)
    proc.send_signal(signal.SIGINT)
    out, err = proc.communicate(timeout=10)
    assert proc.returncode == 0 and out == "" and err == ""


def test_descriptor_table_is_sized_before_threads_exist(hola_server):
    """Once threads share the descriptor table, each doubling of it (at the 65th, 129th, ...
    descriptor) waits for an RCU grace period - tens of milliseconds added to start-up on any
    machine with more than a dozen CPUs, and later to some unlucky accept. The service makes the
    table large while it is still single-threaded; /proc shows the size."""
    srv = hola_server(threads=4)
    status = open("/proc/%d/status" % srv.proc.pid).read()
    fdsize = int([ln for ln in status.splitlines() if ln.startswith("FDSize:")][0].split()[1])
    import resource
    soft, hard = resource.getrlimit(resource.RLIMIT_NOFILE)
    limit = hard if hard != resource.RLIM_INFINITY else 1 << 30  # the service raises its soft limit to the hard one
    assert fdsize >= min(16384, limit - 1), fdsize
