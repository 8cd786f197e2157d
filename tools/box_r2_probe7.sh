#!/usr/bin/env bash
# Round 2, seventh pass: the 32-vCPU shape (8-GPU box), where the round-1 review measured one
# instance with 16 workers falling 28 % per core behind eight small instances. Placement modes at
# (nproc-1)/2 workers, then bench at N = 1 and N = 8. Results in gpurun_out/r2_probe7/.
set -u
OUT=gpurun_out/r2_probe7
mkdir -p "$OUT"
NCPU=$(nproc); echo "nproc=$NCPU" | tee "$OUT/box.txt"
python -c "import __graft_entry__ as g; g.build()" > "$OUT/build.txt" 2>&1
HALF=$(( (NCPU - 1) / 2 ))
python tools/accept_sweep.py --workers $HALF --seconds 6 --repeat 2 --out "$OUT/accept_sweep.json" > "$OUT/accept_sweep.txt" 2>&1
cat "$OUT/accept_sweep.txt"
python bench.py --steps 20 --warmup 3 > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29608 bench.py --gpus 8 --steps 20 --warmup 3 > "$OUT/bench_n8.json" 2> "$OUT/bench_n8.err"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2_probe7/bench_n*.json")):
    try:
        r = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    lat = r.get("latency_us", {})
    print(f.split("/")[-1], "N", r.get("n_gpus"), r["config"].get("io_backend"), "rps", round(r.get("value", 0)), "cores", r.get("server_cores_total"), "per-core", r.get("per_server_core_rps"),
          "spread", r.get("worker_spread"), "p50", lat.get("p50"), "p99", lat.get("p99"), "half", (r.get("half_load") or {}).get("p99_us"), "cold", (r.get("coldstart_ms") or {}).get("median"))
PY
