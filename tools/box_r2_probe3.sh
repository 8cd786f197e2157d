#!/usr/bin/env bash
# Round 2, third pass: the placement question needs a box with at least ~32 vCPUs (15+ workers).
# Prints the CPU count first; when the box is too small for the question it stops there.
# Results in gpurun_out/r2_probe3/.
set -u
OUT=gpurun_out/${OUT_NAME:-r2_probe3}
mkdir -p "$OUT"
NCPU=$(nproc)
{ echo "nproc=$NCPU"; nvidia-smi -L | wc -l; lscpu | grep -E "Model name|Socket|Core|Thread|NUMA|L3"; } > "$OUT/box.txt" 2>&1
cat "$OUT/box.txt"
if [ "$NCPU" -lt "${MIN_CPUS:-30}" ]; then echo "box too small for the 15-worker sweep"; exit 0; fi
python -c "import __graft_entry__ as g; g.build()" > "$OUT/build.txt" 2>&1
HALF=$(( (NCPU - 1) / 2 ))
python tools/accept_sweep.py --workers 4,8,$HALF --seconds ${SWEEP_SECONDS:-6} --repeat ${SWEEP_REPEAT:-3} --out "$OUT/accept_sweep.json" > "$OUT/accept_sweep.txt" 2>&1
cat "$OUT/accept_sweep.txt"
python bench.py --steps 20 --warmup 3 > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 20 --warmup 3 > "$OUT/bench_n$N.json" 2> "$OUT/bench_n$N.err"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/"+__import__("os").environ.get("OUT_NAME","r2_probe3")+"/bench_n*.json")):
    try:
        r = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    lat = r.get("latency_us", {})
    print(f.split("/")[-1], "N", r.get("n_gpus"), "rps", round(r.get("value", 0)), "cores", r.get("server_cores_total"), "per-core", r.get("per_server_core_rps"),
          "spread", r.get("worker_spread"), "p50", lat.get("p50"), "p99", lat.get("p99"), "half", (r.get("half_load") or {}).get("p99_us"))
PY
