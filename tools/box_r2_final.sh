#!/usr/bin/env bash
# Round 2, the record run on the benchmark box with the final code: tests, the headline at 60 s on
# both I/O backends, the N-instance series, cold start, the config-5 proxy, chaining outcomes and
# tracing cost. Results in gpurun_out/r2_final/.
set -u
OUT=gpurun_out/r2_final
mkdir -p "$OUT"
{ nproc; uname -r; cat /proc/sys/kernel/io_uring_disabled; } > "$OUT/box.txt" 2>&1
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > "$OUT/build_smoke.txt" 2>&1; echo "build+smoke rc=$?" >> "$OUT/build_smoke.txt"; tail -2 "$OUT/build_smoke.txt"
build/hola_tests > "$OUT/unit.txt" 2>&1; tail -1 "$OUT/unit.txt"
python -m pytest tests -x -q -m "not gpu" > "$OUT/pytest_not_gpu.txt" 2>&1; tail -2 "$OUT/pytest_not_gpu.txt"
python -m pytest tests -x -q -m gpu > "$OUT/pytest_gpu.txt" 2>&1; tail -2 "$OUT/pytest_gpu.txt"
python bench.py --steps 60 --warmup 3 > "$OUT/bench_n1_60s.json" 2> "$OUT/bench.err"
python bench.py --steps 60 --warmup 3 --io-backend epoll > "$OUT/bench_n1_60s_epoll.json" 2>> "$OUT/bench.err"
python bench.py --steps 20 --warmup 3 --mode conf > "$OUT/bench_n1_conf.json" 2>> "$OUT/bench.err"
python bench.py --steps 20 --warmup 3 --mode notrace > "$OUT/bench_n1_notrace.json" 2>> "$OUT/bench.err"
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 20 --warmup 3 > "$OUT/bench_n$N.json" 2> "$OUT/bench_n$N.err"
done
env -u HOSTNAME -u ZIPKIN_SERVER_URL ./build/hola-coldstart --port 18084 --runs 15 -- ./build/hola --port 18084 --bind 127.0.0.1 --quiet > "$OUT/coldstart_dynamic.json" 2>&1
make static > /dev/null 2>&1
env -u HOSTNAME -u ZIPKIN_SERVER_URL ./build/hola-coldstart --port 18084 --runs 15 -- ./build/hola-static --port 18084 --bind 127.0.0.1 --quiet > "$OUT/coldstart_static.json" 2>&1
env -u HOSTNAME -u ZIPKIN_SERVER_URL ./build/hola-coldstart --port 18084 --runs 15 -- ./build/hola --port 18084 --bind 127.0.0.1 --quiet --io-backend epoll > "$OUT/coldstart_dynamic_epoll.json" 2>&1
python tools/config5_proxy.py --seconds 60 --out "$OUT/config5_proxy.json" > /dev/null 2> "$OUT/config5_proxy.err"
python tools/chaining_paths.py --seconds 8 --out "$OUT/chaining.json" > "$OUT/chaining.txt" 2>&1
python tools/tracing_cost.py --seconds 5 --out "$OUT/tracing.json" > "$OUT/tracing.txt" 2>&1
cat "$OUT/chaining.txt" "$OUT/tracing.txt"
cat "$OUT"/coldstart_*.json
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2_final/bench_n*.json")) + ["gpurun_out/r2_final/config5_proxy.json"]:
    try:
        r = json.loads(open(f).read().strip().splitlines()[-1]) if f.endswith("proxy.json") is False else json.load(open(f))
    except Exception as e:
        print(f, "unreadable", e); continue
    if "value" not in r:
        print(f.split("/")[-1], "rps", round(r["rps"]), r["latency_us"], "errors", r["errors"], r["startup_line"]); continue
    lat = r.get("latency_us", {}); h = r.get("half_load") or {}
    print(f.split("/")[-1], "N", r.get("n_gpus"), r["config"].get("io_backend"), "rps", round(r["value"]), "cores", r.get("server_cores_total"), "per-core", r.get("per_server_core_rps"),
          "spread", r.get("worker_spread"), "p50", lat.get("p50"), "p99", lat.get("p99"), "p999", lat.get("p999"), "half", h.get("p50_us"), h.get("p99_us"), h.get("p999_us"),
          "cold", (r.get("coldstart_ms") or {}).get("median"), "cpu", r["config"].get("server_cpu_us_per_request"))
PY
