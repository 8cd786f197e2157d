"""The C++ unit tests, the load generator and the bench/driver entry points."""
from __future__ import annotations

import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

from hola_b200 import client, loadgen

REPO = Path(__file__).resolve().parent.parent


def test_cpp_unit_tests_pass(native_build):
    out = subprocess.run([str(native_build.unit_tests)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "0 failed" in out.stdout


def test_loadgen_closed_loop_counts_and_verifies(hola_server):
    srv = hola_server()
    rep = loadgen.run(srv.base_url + "/api/hola", connections=8, duration=1.0, warmup=0.5, interval=0.5, threads=2,
                      expect_body=b"Hola de unknown")
    assert rep["errors"]["total"] == 0 and rep["mode"] == "closed"
    assert rep["requests"] > 100 and rep["rps"] > 100
    assert len(rep["intervals"]) == 3 and rep["intervals"][0]["warmup"] is True and rep["intervals"][1]["warmup"] is False
    assert rep["requests"] == sum(i["requests"] for i in rep["intervals"] if not i["warmup"])
    lat = rep["latency_us"]
    assert 0 < lat["min"] <= lat["p50"] <= lat["p99"] <= lat["max"]
    # The server agrees on the order of magnitude of work done.
    assert rep["bytes"] > rep["requests"] * 100


def test_loadgen_detects_wrong_bodies_and_statuses(hola_server):
    srv = hola_server()
    rep = loadgen.run(srv.base_url + "/api/hola", connections=2, duration=0.3, warmup=0, interval=0.3, expect_body=b"something else",
                      allow_errors=True)
    assert rep["errors"]["body"] > 0 and rep["requests"] == 0
    rep = loadgen.run(srv.base_url + "/api/missing", connections=2, duration=0.3, warmup=0, interval=0.3, allow_errors=True)
    assert rep["errors"]["status"] > 0
    with pytest.raises(RuntimeError, match="errors"):
        loadgen.run(srv.base_url + "/api/missing", connections=1, duration=0.2, warmup=0, interval=0.2)
    rep = loadgen.run(srv.base_url + "/api/missing", connections=2, duration=0.3, warmup=0, interval=0.3, expect_status=404)
    assert rep["errors"]["total"] == 0 and rep["requests"] > 0


def test_loadgen_open_loop_holds_the_rate(hola_server):
    srv = hola_server()
    rep = loadgen.run(srv.base_url + "/api/hola", connections=4, duration=2.0, warmup=0.5, interval=0.5, rate=2000, threads=2)
    assert rep["mode"] == "open" and rep["errors"]["total"] == 0
    assert 0.9 * 2000 <= rep["rps"] <= 1.1 * 2000, rep["rps"]


def test_loadgen_pipelining_http10_and_connection_refused(hola_server):
    srv = hola_server()
    rep = loadgen.run(srv.base_url + "/api/hola", connections=2, duration=0.5, warmup=0, interval=0.5, pipeline=8,
                      expect_body=b"Hola de unknown")
    assert rep["errors"]["total"] == 0 and rep["requests"] > 100
    rep = loadgen.run(srv.base_url + "/api/health", connections=2, duration=0.3, warmup=0, interval=0.3, http10=True,
                      expect_body=b"I'm ok")
    assert rep["errors"]["total"] == 0 and rep["reconnects"] == 0  # HTTP/1.0 keep-alive honoured
    rep = loadgen.run("http://127.0.0.1:9/", connections=2, duration=0.3, warmup=0, interval=0.3, allow_errors=True)
    assert rep["errors"]["connect"] > 0 and rep["requests"] == 0


def test_loadgen_against_chunked_and_closing_servers(fake_aloha):
    fake_aloha.mode = "chunked"
    rep = loadgen.run(fake_aloha.url + "api/aloha-chaining", connections=2, duration=0.5, warmup=0, interval=0.5,
                      expect_body=json.dumps(fake_aloha.payload).encode())
    assert rep["errors"]["total"] == 0 and rep["requests"] > 5


def _run_bench(*args, env=None, timeout=180):
    e = dict(os.environ)
    e.update(env or {})
    out = subprocess.run([sys.executable, str(REPO / "bench.py"), *args], capture_output=True, text=True, timeout=timeout, env=e)
    assert out.returncode == 0, out.stderr[-3000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, out.stdout
    return json.loads(lines[0])


@pytest.mark.slow
def test_bench_prints_one_contract_line():
    r = _run_bench("--steps", "2", "--warmup", "3")
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches"):
        assert key in r, key
    assert r["n_gpus"] == 1 and r["steps"] == 2 and r["warmup"] == 3
    assert r["value"] > 1000 and r["unit"] == "req/s" and r["higher_is_better"] is True
    assert r["gpu_launches"] == 0  # stated, not hidden: there is no GPU work in this project
    assert r["vs_baseline"] is None  # the reference publishes no number
    assert r["e2e"]["h2d_bytes_per_step"] == 0 and r["e2e"]["value"] == r["value"]
    assert r["config"]["connections_per_instance"] == 64
    assert r["latency_us"]["p50"] <= r["latency_us"]["p99"]
    assert "GET /api/hola" in r["metric"]
    # The keys that make the line readable on its own.
    assert r["server_cores_total"] == len(r["config"]["server_cpus"]) >= 1
    assert abs(r["per_server_core_rps"] * r["server_cores_total"] - r["value"]) < r["server_cores_total"]
    assert r["config"]["accept_mode"] == "balanced" and 1.0 <= r["worker_spread"] < 2.0
    assert sum(r["config"]["worker_connections"]) == 64
    assert max(r["config"]["worker_connections"]) - min(r["config"]["worker_connections"]) <= 1
    assert r["spans"]["dropped"] > 0 and r["spans"]["reported"] == 0  # as-is: every span goes to a collector nobody runs
    assert 0.05 < r["coldstart_ms"]["median"] < 2000 and r["coldstart_ms"]["runs"] == 5
    assert "cold_start_s" not in r["config"] and "listen_log_s" in r["config"]
    hk = r["config"]["housekeeping_cpu"]
    if len(os.sched_getaffinity(0)) >= 4:
        assert hk is not None and hk not in r["config"]["server_cpus"] and hk not in r["config"]["load_cpus"]


def test_bench_reference_arm_reports_unavailable():
    env = {k: v for k, v in os.environ.items() if k != "HOLA_REFERENCE_CMD"}
    out = subprocess.run([sys.executable, str(REPO / "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, timeout=60, env=env)
    assert out.returncode == 0
    r = json.loads(out.stdout.strip().splitlines()[-1])
    assert r["impl"] == "reference" and "unavailable" in r and len(r["unavailable"]) > 20


@pytest.mark.slow
def test_bench_reference_arm_drives_another_program_through_the_same_procedure():
    """No box has a JVM, so the reference arm has never met the real jar. What can be shown is
    that the arm is wired: given a launch command (HOLA_REFERENCE_CMD, or java + the jar under
    baseline/_ref) it starts that program pinned to the server CPUs, waits for its first 200, runs
    the identical hola-load invocation and prints the same schema - here against a stand-in that
    is not this repository's server. Two such lines are all the driver needs for a ratio."""
    standin = "%s %s {port}" % (sys.executable, REPO / "tests" / "data" / "standin_reference.py")
    ref = _run_bench("--impl", "reference", "--steps", "2", "--warmup", "3", "--coldstart-runs", "2", "--half-load-seconds", "0",
                     env={"HOLA_REFERENCE_CMD": standin})
    ours = _run_bench("--steps", "2", "--warmup", "3", "--coldstart-runs", "2", "--half-load-seconds", "0")
    assert ref["impl"] == "reference" and ours["impl"] == "ours"
    assert set(ref) == set(ours) and set(ref["config"]) == set(ours["config"])
    for key in ("metric", "unit", "n_gpus", "steps", "warmup", "data"):
        assert ref[key] == ours[key]
    assert ref["config"]["server_cpus"] == ours["config"]["server_cpus"] and ref["config"]["load_cpus"] == ours["config"]["load_cpus"]
    assert ref["value"] > 100 and ref["requests"] > 200 and ref["errors"] == 0
    assert ref["coldstart_ms"]["median"] > ours["coldstart_ms"]["median"]  # an interpreter starts slower than the executable
    assert ref["worker_spread"] is None and ref["spans"] is None          # nothing the reference can be asked for
    assert ours["value"] / ref["value"] > 1.0                             # the ratio exists (and a Python stand-in is slow)


@pytest.mark.slow
def test_bench_two_instances_in_one_process():
    r = _run_bench("--gpus", "2", "--steps", "1", "--warmup", "3")
    assert r["n_gpus"] == 2 and r["config"]["global_batch"] == 128 and r["value"] > 1000


@pytest.mark.slow
def test_bench_under_torchrun_two_ranks():
    e = dict(os.environ)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                          "127.0.0.1", "--master-port", "29617", str(REPO / "bench.py"), "--gpus", "2", "--steps", "1", "--warmup", "3"],
                         capture_output=True, text=True, timeout=300, env=e)
    assert out.returncode == 0, out.stderr[-3000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    r = json.loads(lines[0])
    assert r["n_gpus"] == 2 and r["requests"] > 1000


def test_differential_tool_same_and_different(hola_server):
    """tools/differential.py: the harness kept ready for a runnable Java reference."""
    a = hola_server()
    b = hola_server(args=["--accept-mode", "rr"], threads=1)
    c = hola_server(args=["-Dhello=Bonjour de %s"])
    tool = str(REPO / "tools" / "differential.py")
    same = subprocess.run([sys.executable, tool, a.base_url, b.base_url, "--json"], capture_output=True, text=True, timeout=120)
    assert same.returncode == 0, same.stdout[-2000:] + same.stderr[-2000:]
    rep = json.loads(same.stdout)
    assert rep["requests"] >= 20 and rep["contract_differences"] == 0
    diff = subprocess.run([sys.executable, tool, a.base_url, c.base_url, "--json"], capture_output=True, text=True, timeout=120)
    assert diff.returncode == 1
    rep = json.loads(diff.stdout)
    changed = [r["request"] for r in rep["report"] if r["contract"]]
    assert any("/api/hola" in r for r in changed) and not any("/api/health" in r for r in changed)


@pytest.mark.slow
def test_differential_tool_reports_every_unverified_decision(native_build):
    """DESIGN.md section 4 lists what was decided from memory of libraries that cannot be run here.
    The differential tool turns that list into a machine-readable report: when it starts both
    services itself (so that it owns their span collectors, their logs and their greeting pattern)
    every item is compared; here both sides are this service, so every compared item is `same`."""
    tool = str(REPO / "tools" / "differential.py")
    hola = "%s --port {port} --bind 127.0.0.1 --zipkin-flush-ms 50 --aloha-url http://127.0.0.1:9/ --feign-max-attempts 1" % native_build.hola
    out = subprocess.run([sys.executable, tool, "--launch-left", hola + " --threads 1 --reference-logging",
                          "--launch-right", hola + " --threads 2 --accept-mode rr --log-feign", "--json"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    rep = json.loads(out.stdout)
    items = {i["id"]: i for i in rep["items"]}
    compared = {"response-head", "error-bodies", "chaining-json-whitespace", "swagger-text", "hystrix-stream-fields", "zipkin-span-json",
                "b3-ids-need-sampled-header", "feign-basic-log", "formatter-corner-cases"}
    assert compared <= set(items)
    assert {i for i in compared if items[i]["status"] != "same"} == set(), {i: items[i] for i in compared if items[i]["status"] != "same"}
    assert items["feign-basic-log"]["details"]["left"]  # the lines were really there to compare
    assert items["dns-cache-ttl"]["status"] == "not-compared" and items["feign-retryer-schedule"]["status"] == "not-compared"
    # Against something else the same items come back `different` (or not comparable), not as a crash.
    standin = "%s %s {port}" % (sys.executable, REPO / "tests" / "data" / "standin_reference.py")
    out = subprocess.run([sys.executable, tool, "--launch-left", hola + " --threads 1", "--launch-right", standin, "--json"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 1
    items = {i["id"]: i for i in json.loads(out.stdout)["items"]}
    assert items["response-head"]["status"] == "different" and items["swagger-text"]["status"] == "different"
    assert items["formatter-corner-cases"]["status"] == "not-compared"


def test_graft_entry_build_and_smoke():
    sys.path.insert(0, str(REPO))
    import __graft_entry__
    __graft_entry__.build()
    __graft_entry__.smoke()
