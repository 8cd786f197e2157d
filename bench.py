#!/usr/bin/env python
"""Headline benchmark: GET /api/hola requests/sec and p50/p99 latency, CPU, localhost.

That is the metric BASELINE.json names for objectiser/hola, on its config 4: 64 concurrent
keep-alive clients, sustained load. One *step* is one second of that load; ``--steps K`` measures
K seconds after ``--warmup W`` seconds, so the defaults (K=60, W=3) are exactly the 60 s run the
baseline describes.

    python bench.py [--gpus N] [--steps K] [--warmup W]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference      # same procedure against the Java service, when one can be launched

What "N GPUs" means for a service that never touches a GPU: N independent service instances
(the reference scales out as "more pods behind a LoadBalancer", /root/reference/pom.xml:80), each
with its own 64-connection load generator, on disjoint slices of the machine's CPUs. Under
torchrun every rank runs one instance and the totals are summed over ranks (gloo); without
torchrun one process runs all N. ``value`` is the whole-job aggregate requests/sec.

CPU plan. One CPU (the highest-numbered one) is kept out of every slice as *housekeeping*: this
script, its clock sampler and whatever else the box runs meanwhile live there, so that nothing has
to preempt a pinned server or generator thread to get CPU time. The rest is cut into N contiguous
slices; in each, the first half (rounded down) serves and the second half generates load.
Because every N divides the same CPUs, aggregate throughput cannot grow with N: the figure to
compare across N is ``per_server_core_rps``.

The line printed at the end is honest about the GPU: ``gpu_launches`` is 0 because the workload
has no kernel to launch (BASELINE.json north_star). The measurement is end to end
by construction - real TCP connections, every response parsed and its body verified byte for
byte - so ``e2e`` is the same number, not a second one.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time
from pathlib import Path
from typing import Dict, List, Optional, Sequence, Tuple

REPO = Path(__file__).resolve().parent
sys.path.insert(0, str(REPO))

METRIC = "GET /api/hola requests/sec (p50/p99 latency alongside), CPU, localhost"
UNIT = "req/s"
CONNECTIONS = 64
ZIPKIN_DEAD_URL = "http://127.0.0.1:9"  # discard port: nothing listens, like zipkin-query:9411 on a dev box


def partition(cpus: Sequence[int], parts: int) -> List[List[int]]:
    """Contiguous, near-equal slices; with fewer CPUs than parts the slices overlap round-robin."""
    cpus = list(cpus)
    if len(cpus) >= parts:
        base, extra = divmod(len(cpus), parts)
        out, pos = [], 0
        for i in range(parts):
            n = base + (1 if i < extra else 0)
            out.append(cpus[pos:pos + n])
            pos += n
        return out
    return [[cpus[i % len(cpus)]] for i in range(parts)]


def plan_cpus(cpus: Sequence[int], parts: int) -> Tuple[Optional[int], List[List[int]]]:
    """(housekeeping CPU or None, one slice per instance). A CPU is set aside only when the
    machine can spare it: at least 4 CPUs, and every instance still gets a CPU for the server and
    one for the generator."""
    cpus = list(cpus)
    if len(cpus) >= 4 and len(cpus) - 1 >= 2 * parts:
        return cpus[-1], partition(cpus[:-1], parts)
    return None, partition(cpus, parts)


def mode_settings(mode: str):
    """(env, extra args, expected body) for one of the three configurations."""
    env: Dict[str, Optional[str]] = {"HOSTNAME": None}  # a plain `mvn wildfly-swarm:run` from a shell does not export HOSTNAME
    args: List[str] = []
    expect = b"Hola de unknown"
    if mode == "as-is":
        # The reference cannot even deploy without ZIPKIN_SERVER_URL, so the
        # as-is configuration traces every request and ships spans to a collector nobody runs.
        env["ZIPKIN_SERVER_URL"] = ZIPKIN_DEAD_URL
    elif mode == "notrace":
        env["ZIPKIN_SERVER_URL"] = None
    elif mode == "conf":
        env["ZIPKIN_SERVER_URL"] = ZIPKIN_DEAD_URL
        args += ["-Dconf=" + str(REPO / "tests" / "data" / "translation.properties")]
        expect = b"Hola de unknown mediante el archivo de configuraciones\n"
    return env, args, expect


def make_server(impl: str, mode: str, server_cpus: Sequence[int], threads: int, io_backend: str = "auto"):
    """The process under test: this repository's executable, or the reference (hola_b200/reference.py)."""
    env, args, _ = mode_settings(mode)
    if io_backend != "auto":
        args = args + ["--io-backend", io_backend]
    if impl == "reference":
        from hola_b200.reference import ReferenceServer, reference_command
        command, _why = reference_command()
        if mode == "conf":
            command = command[:1] + args + command[1:]  # -Dconf=... right after `java`
        if env.get("ZIPKIN_SERVER_URL") is None:
            env["ZIPKIN_SERVER_URL"] = ZIPKIN_DEAD_URL  # it cannot start without one
        return ReferenceServer(command, cpus=server_cpus, env=env)
    from hola_b200.server import HolaServer
    return HolaServer(args=["--cpus", ",".join(map(str, server_cpus))] + args, env=env, threads=threads)


def coldstart_ms(impl: str, mode: str, runs: int = 5) -> Optional[Dict]:
    """BASELINE.json config 3: exec -> first HTTP 200, median of ``runs`` launches (hola-coldstart)."""
    from hola_b200 import loadgen, paths
    from hola_b200.reference import free_port, reference_command
    env_over, args, _ = mode_settings(mode)
    env = dict(os.environ)
    for k, v in env_over.items():
        if v is None:
            env.pop(k, None)
        else:
            env[k] = v
    port = free_port()
    if impl == "reference":
        command, _why = reference_command()
        command = [part.replace("{port}", str(port)) for part in command]
        env.setdefault("ZIPKIN_SERVER_URL", ZIPKIN_DEAD_URL)
        timeout_ms = 180000
    else:
        command = [str(paths().hola), "--port", str(port), "--bind", "127.0.0.1", "--quiet"] + args
        timeout_ms = 10000
    try:
        rep = loadgen.cold_start(command, port, runs=runs, timeout_ms=timeout_ms, env=env)
        return {"median": round(rep["median"] / 1000.0, 3), "min": round(rep["min"] / 1000.0, 3), "max": round(rep["max"] / 1000.0, 3),
                "runs": runs, "what": "exec -> first HTTP 200 on /api/hola, default thread count"}
    except Exception as exc:  # noqa: BLE001 - additional information only
        return {"error": "%s: %s" % (type(exc).__name__, exc)}


def run_instance(index: int, cpus: Sequence[int], steps: int, warmup: int, mode: str, impl: str, results: Dict[int, Dict],
                 half_load_seconds: int = 0, io_backend: str = "auto") -> None:
    from hola_b200 import loadgen

    cpus = list(cpus)
    if len(cpus) >= 2:
        n_server = len(cpus) // 2  # an odd CPU goes to the generator: the server must be the side that saturates
        server_cpus, load_cpus = cpus[:n_server], cpus[n_server:]
    else:
        server_cpus = load_cpus = cpus
    server_threads = min(len(server_cpus), CONNECTIONS)
    load_threads = min(len(load_cpus), CONNECTIONS)
    _, _, expect = mode_settings(mode)
    with make_server(impl, mode, server_cpus, server_threads, io_backend) as srv:
        before = srv.stats()
        report = loadgen.run(srv.base_url + "/api/hola", connections=CONNECTIONS, duration=steps, warmup=warmup,
                             interval=1.0, threads=load_threads, cpus=load_cpus, expect_body=expect)
        after = srv.stats()
        report["server_cpus"] = list(server_cpus)
        report["load_cpus"] = list(load_cpus)
        report["server_threads"] = server_threads
        report["startup_s"] = srv.startup_seconds
        if before and after:
            per_worker = [a["requests"] - b["requests"] for a, b in zip(after["workers"], before["workers"])]
            total = max(1, after["requests"] - before["requests"])
            report["worker_spread"] = round(max(per_worker) / max(1, min(per_worker)), 3)
            report["worker_connections"] = [a["connections"] - b["connections"] for a, b in zip(after["workers"], before["workers"])]
            report["spans"] = {k: after["spans"][k] - before["spans"][k] for k in ("reported", "dropped")}
            report["server_cpu_us_per_request"] = {
                "user": round((after["cpu_user_s"] - before["cpu_user_s"]) * 1e6 / total, 3),
                "sys": round((after["cpu_sys_s"] - before["cpu_sys_s"]) * 1e6 / total, 3)}
            report["accept_mode"] = after["accept_mode"]
            report["io_backend"] = after.get("io_backend")
        if half_load_seconds > 0:
            # Context for the headline, outside its timed region: the same 64 connections driven
            # open-loop at half the measured throughput, latency taken from the scheduled send time.
            # This is what a client sees when the service is busy but not saturated.
            try:
                half = loadgen.run(srv.base_url + "/api/hola", connections=CONNECTIONS, duration=half_load_seconds, warmup=1,
                                   interval=1.0, threads=load_threads, cpus=load_cpus, expect_body=expect, rate=report["rps"] / 2)
                report["half_load"] = {"offered_rps": round(report["rps"] / 2, 1), "achieved_rps": half["rps"],
                                       "p50_us": half["latency_us"]["p50"], "p99_us": half["latency_us"]["p99"],
                                       "p999_us": half["latency_us"]["p999"], "seconds": half_load_seconds}
            except Exception as exc:  # noqa: BLE001 - additional information only
                report["half_load"] = {"error": "%s: %s" % (type(exc).__name__, exc)}
    results[index] = report


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1, help="number of service instances (see module docstring)")
    ap.add_argument("--steps", type=int, default=60, help="measured seconds (one step = one second of load)")
    ap.add_argument("--warmup", type=int, default=3, help="warm-up seconds, not measured (>= 3)")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--mode", choices=["as-is", "notrace", "conf"], default="as-is",
                    help="as-is: tracing on, collector unreachable (how the reference must be run); "
                         "notrace: ZIPKIN_SERVER_URL unset; conf: greeting from translation.properties")
    ap.add_argument("--max-cpus", type=int, default=int(os.environ.get("HOLA_BENCH_MAX_CPUS", "0")),
                    help="use at most this many CPUs in total (0: all the process may run on)")
    ap.add_argument("--half-load-seconds", type=int, default=5,
                    help="after the timed run (N=1 only): an open-loop run at half the measured rate for this many "
                         "seconds, reported as `half_load` (0 to skip)")
    ap.add_argument("--coldstart-runs", type=int, default=5, help="launches for `coldstart_ms` (0 to skip)")
    ap.add_argument("--io-backend", choices=["auto", "epoll", "uring"], default=os.environ.get("HOLA_BENCH_IO_BACKEND", "auto"),
                    help="the service's --io-backend (ours only; auto is the service's default: io_uring where the kernel "
                         "offers it, else epoll - what was used is reported as config.io_backend)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3  # timing rule: W >= 3

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    distributed = world > 1
    n = args.gpus if args.gpus > 0 else 1

    if args.impl == "reference":
        from hola_b200.reference import reference_command
        command, why = reference_command()
        if command is None:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": why, "metric": METRIC, "n_gpus": n}))
            return 0

    import __graft_entry__
    if int(os.environ.get("LOCAL_RANK", "0")) == 0:
        __graft_entry__.build()  # the load generator is needed for either arm

    dist = None
    if distributed:
        import torch
        import torch.distributed as dist  # noqa: F811
        # gloo on purpose: the only thing exchanged is a handful of counters held in host memory.
        dist.init_process_group(backend="gloo")
        dist.barrier()  # everyone waits for rank 0's build

    from hola_b200.utils import ClockSampler, available_cpus, cpu_info

    cpus = available_cpus()
    if args.max_cpus > 0:
        cpus = cpus[:args.max_cpus]
    housekeeping, slices = plan_cpus(cpus, n)
    # Cold start first, while this process may still run anywhere: the launched service sizes
    # itself from the CPUs it is allowed to use, as it would when started by hand.
    cold = coldstart_ms(args.impl, args.mode, args.coldstart_runs) if (rank == 0 and args.coldstart_runs > 0) else None
    if housekeeping is not None:
        # From here on this process, its threads and every child that does not pin itself (the
        # clock sampler's nvidia-smi, the service's reporter and resolver threads) stay off the
        # measured CPUs.
        os.sched_setaffinity(0, {housekeeping})
    my_indices = [rank] if distributed else list(range(n))
    results: Dict[int, Dict] = {}

    sampler = ClockSampler().start() if rank == 0 else None
    if dist is not None:
        dist.barrier()
    t0 = time.monotonic()

    def guarded(i: int) -> None:
        try:
            run_instance(i, slices[i % len(slices)], args.steps, args.warmup, args.mode, args.impl, results,
                         half_load_seconds=args.half_load_seconds if (n == 1 and not distributed) else 0, io_backend=args.io_backend)
        except Exception as exc:  # noqa: BLE001 - reported below, on every rank
            results[i] = {"error": "%s: %s" % (type(exc).__name__, exc)}

    threads = [threading.Thread(target=guarded, args=(i,)) for i in my_indices]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    wall = time.monotonic() - t0
    if dist is not None:
        dist.barrier()
    clocks = sampler.stop() if sampler else None

    # A failed instance (a request error, a server that did not start) invalidates the whole run.
    # Every rank learns about it before anyone leaves, so that nobody is left waiting in a collective.
    missing = [i for i in my_indices if i not in results or "error" in results[i]]
    failed = 1.0 if missing else 0.0
    if dist is not None:
        import torch
        flag = torch.tensor([failed], dtype=torch.float64)
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        failed = float(flag[0])
    if failed:
        for i in missing:
            print("bench: instance %d failed: %s" % (i, results.get(i, {}).get("error", "no result")), file=sys.stderr)
        if dist is not None:
            dist.destroy_process_group()
        return 1

    # Aggregate over local instances, then over ranks.
    local_requests = sum(r["requests"] for r in results.values())
    local_rps = sum(r["rps"] for r in results.values())
    local_bytes = sum(r["bytes"] for r in results.values())
    server_cores = float(sum(len(r["server_cpus"]) for r in results.values()))
    spans_reported = float(sum((r.get("spans") or {}).get("reported", 0) for r in results.values()))
    spans_dropped = float(sum((r.get("spans") or {}).get("dropped", 0) for r in results.values()))
    p50 = max(r["latency_us"]["p50"] for r in results.values())
    p99 = max(r["latency_us"]["p99"] for r in results.values())
    p999 = max(r["latency_us"]["p999"] for r in results.values())
    lat_max = max(r["latency_us"]["max"] for r in results.values())
    spread = max((r.get("worker_spread") or 0.0) for r in results.values())
    step_rps = [sum(r["intervals"][k]["rps"] for r in results.values()) for k in range(args.warmup, args.warmup + args.steps)]
    if dist is not None:
        import torch
        sums = torch.tensor([float(local_requests), float(local_rps), float(local_bytes), server_cores, spans_reported, spans_dropped],
                            dtype=torch.float64)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
        maxes = torch.tensor([p50, p99, p999, lat_max, wall, spread], dtype=torch.float64)
        dist.all_reduce(maxes, op=dist.ReduceOp.MAX)  # latency, spread: the worst instance; time: the slowest rank
        steps_t = torch.tensor(step_rps, dtype=torch.float64)
        dist.all_reduce(steps_t, op=dist.ReduceOp.SUM)
        local_requests, local_rps, local_bytes, server_cores, spans_reported, spans_dropped = (float(x) for x in sums)
        p50, p99, p999, lat_max, wall, spread = (float(x) for x in maxes)
        step_rps = [float(x) for x in steps_t]
        dist.destroy_process_group()
    if rank != 0:
        return 0

    first = results[my_indices[0]]
    baseline = json.loads((REPO / "BASELINE.json").read_text())
    # value / published number, should BASELINE.json ever carry one for requests/sec.
    published = (baseline.get("published") or {}).get("requests_per_sec")
    vs_baseline = round(local_rps / float(published), 3) if published else None
    ours = args.impl == "ours"
    out = {
        "metric": METRIC,
        "value": round(local_rps, 1),
        "unit": UNIT,
        "n_gpus": n,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": 1000.0,  # a step is defined as one second of sustained load
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": vs_baseline,  # null: the reference publishes no number
        "dtype": "n/a (no tensor math; bytes over TCP)",
        "data": "synthetic: 64 keep-alive HTTP/1.1 clients on localhost, every body verified byte for byte",
        "impl": args.impl,
        "latency_us": {"p50": p50, "p99": p99, "p999": p999, "max": lat_max},
        "requests": int(local_requests),
        "errors": 0,
        "step_rps": {"min": round(min(step_rps), 1), "median": round(statistics.median(step_rps), 1), "max": round(max(step_rps), 1)},
        # The three keys that make runs on different boxes and at different N comparable.
        "server_cores_total": int(server_cores),
        "per_server_core_rps": round(local_rps / max(1.0, server_cores), 1),
        "worker_spread": round(spread, 3) if spread else None,  # max/min requests per server thread (worst instance); 1.0 = even
        "spans": {"reported": int(spans_reported), "dropped": int(spans_dropped)} if ours else None,
        "coldstart_ms": cold,
        "config": {
            "model": "objectiser/hola GET /api/hola (BASELINE.json config 4%s)" % ("" if args.steps == 60 else "; %d s instead of its 60 s" % args.steps),
            "global_batch": CONNECTIONS * n,
            "seq_len": None,
            "parallelism": "%d independent instance(s), one per rank" % n,
            "scaling_note": "every N divides the same %d CPUs (one more is kept free for housekeeping: this script, samplers, the OS), "
                            "half of each slice serving and half generating load, so the driver's efficiency v_N/(N*v_1) is bounded by "
                            "1/N by construction; compare per_server_core_rps across N instead. GPUs are not involved" % (len(cpus) - (housekeeping is not None)),
            "connections_per_instance": CONNECTIONS,
            "mode": args.mode,
            "accept_mode": first.get("accept_mode"),
            "io_backend": first.get("io_backend"),
            "housekeeping_cpu": housekeeping,
            "server_cpus": first["server_cpus"],
            "load_cpus": first["load_cpus"],
            "server_threads": first["server_threads"],
            "worker_connections": first.get("worker_connections"),
            "server_cpu_us_per_request": first.get("server_cpu_us_per_request"),
            "cpu": cpu_info(),
            "l2_flush": "n/a: no device buffers; each request is a fresh socket read/write",
            "wall_s": round(wall, 2),
            "listen_log_s": round(first["startup_s"], 4),  # Popen -> "listening" log line seen by this script; see coldstart_ms for exec -> first 200
        },
        "clocks": {k: clocks[k] for k in ("sm_mhz", "sm_max_mhz", "reasons")} if clocks else None,
        "clocks_note": clocks["note"] if clocks else None,
        "cpu_mhz_median": clocks["cpu_mhz_median"] if clocks else None,
        "e2e": {
            "value": round(local_rps, 1),
            "unit": UNIT,
            "h2d_bytes_per_step": 0,
            "d2h_bytes_per_step": 0,
            "note": "the only measurement there is: requests go through real TCP sockets to the running service and every "
                    "response is read and checked; no device is involved, so there is no device-timed number to contrast it with",
            "wire_bytes_per_step": int(local_bytes / max(1, args.steps)),
        },
        "half_load": first.get("half_load"),
        "gpu_launches": 0,
        "gpu_note": "CPU-only service: the reference has no kernels, tensors or collectives (SURVEY.md 2.3); none were invented",
    }
    print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
