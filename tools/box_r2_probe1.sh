#!/usr/bin/env bash
# Round 2, first pass on the benchmark box: does everything build and pass there, what does the
# new default placement do at the worker counts the box allows, and what does the new bench line
# say (housekeeping CPU, per-core throughput, spread, spans, cold start, half-load latency).
# Results in gpurun_out/r2_probe1/.
set -u
OUT=gpurun_out/r2_probe1
mkdir -p "$OUT"
{ nproc; lscpu | head -25; for t in java javac mvn docker podman; do command -v $t || echo "$t: absent"; done; uname -r; } > "$OUT/box.txt" 2>&1
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > "$OUT/build_smoke.txt" 2>&1; echo "build+smoke rc=$?" >> "$OUT/build_smoke.txt"
build/hola_tests > "$OUT/unit.txt" 2>&1; tail -1 "$OUT/unit.txt"
python -m pytest tests -x -q -m "not gpu" > "$OUT/pytest_not_gpu.txt" 2>&1; tail -3 "$OUT/pytest_not_gpu.txt"
python -m pytest tests -x -q -m gpu > "$OUT/pytest_gpu.txt" 2>&1; tail -3 "$OUT/pytest_gpu.txt"

NCPU=$(nproc)
HALF=$(( (NCPU - 1) / 2 ))
python tools/accept_sweep.py --workers 4,$HALF --seconds 6 --repeat 3 --out "$OUT/accept_sweep.json" > "$OUT/accept_sweep.txt" 2>&1
cat "$OUT/accept_sweep.txt"

python bench.py --steps 20 --warmup 3 > "$OUT/bench_n1_20s.json" 2> "$OUT/bench_n1_20s.err"
python bench.py --steps 60 --warmup 3 > "$OUT/bench_n1_60s.json" 2> "$OUT/bench_n1_60s.err"
python bench.py --impl reference --steps 2 --warmup 3 > "$OUT/bench_reference.json" 2>&1
for N in 2 4; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 10 --warmup 3 > "$OUT/bench_n$N.json" 2> "$OUT/bench_n$N.err"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2_probe1/bench_n*.json")):
    try:
        r = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    lat = r.get("latency_us", {})
    print(f.split("/")[-1], "N", r.get("n_gpus"), "rps", round(r.get("value", 0)), "cores", r.get("server_cores_total"), "per-core", r.get("per_server_core_rps"),
          "spread", r.get("worker_spread"), "p50", lat.get("p50"), "p99", lat.get("p99"), "half", (r.get("half_load") or {}).get("p99_us"),
          "cold", (r.get("coldstart_ms") or {}).get("median"), "spans", r.get("spans"))
PY
