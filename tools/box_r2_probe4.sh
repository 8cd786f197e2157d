#!/usr/bin/env bash
# Round 2, fourth pass: the io_uring backend of the real server against its epoll backend and
# against both null servers, plus the black-box tests on the box's kernel with --io-backend uring.
# Results in gpurun_out/r2_probe4/.
set -u
OUT=gpurun_out/r2_probe4
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build()" > "$OUT/build.txt" 2>&1; echo "build rc=$?" >> "$OUT/build.txt"
python -m pytest tests/test_contract.py tests/test_server_core.py tests/test_chaining.py tests/test_tracing.py tests/test_stream.py tests/test_fuzz.py \
    --io-backend uring -q -x > "$OUT/pytest_uring.txt" 2>&1; tail -3 "$OUT/pytest_uring.txt"
python tools/ceiling.py --seconds 8 --repeat 3 --out "$OUT/ceiling.json" > "$OUT/ceiling.txt" 2>&1
cat "$OUT/ceiling.txt"
for be in epoll uring epoll uring; do
  python bench.py --steps 20 --warmup 3 --io-backend $be --coldstart-runs 0 > "$OUT/bench_${be}_$RANDOM.json" 2>> "$OUT/bench.err"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2_probe4/bench_*.json")):
    try:
        r = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    lat = r.get("latency_us", {}); h = r.get("half_load") or {}
    print(r["config"]["io_backend"], "rps", round(r["value"]), "per-core", r["per_server_core_rps"], "p50", lat.get("p50"), "p99", lat.get("p99"), "p999", lat.get("p999"),
          "half p50/p99/p999", h.get("p50_us"), h.get("p99_us"), h.get("p999_us"), "cpu", r["config"]["server_cpu_us_per_request"])
PY
