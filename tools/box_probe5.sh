#!/usr/bin/env bash
# Fifth pass: does busy-polling (--spin-us) tame the latency tail below saturation? Open loop at
# 1.1 M/s (about half of capacity), 20 s each, four combinations. Results in gpurun_out/probe5/.
set -u
OUT=gpurun_out/probe5
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build()" > "$OUT/build.txt" 2>&1
run() {  # run <label> <server spin> <load spin>
  env -u HOSTNAME ZIPKIN_SERVER_URL=http://127.0.0.1:9 ./build/hola --port 18080 --bind 127.0.0.1 --threads 8 --cpus 0-7 --spin-us "$2" --verbose 2> "$OUT/server_$1.err" &
  local pid=$!
  sleep 0.3
  ./build/hola-load --json -c 64 -d 20 -w 2 -t 8 --cpus 8-15 --rate 1100000 --spin-us "$3" --expect-body "Hola de unknown" http://127.0.0.1:18080/api/hola > "$OUT/$1.json" 2>> "$OUT/server_$1.err"
  kill $pid; wait $pid 2>/dev/null
  grep "stats:" "$OUT/server_$1.err" | sed 's/.*stats: //' > "$OUT/profile_$1.txt"
}
run sleep_sleep 0 0
run spin_sleep 100 0
run sleep_spin 0 100
run spin_spin 100 100
run sleep_sleep_again 0 0
python - <<'EOF'
import json, glob
for f in sorted(glob.glob("gpurun_out/probe5/*.json")):
    r = json.loads(open(f).read().strip().splitlines()[-1])
    lat = r["latency_us"]
    print("%-24s rps=%8d p50=%6.1f p90=%6.1f p99=%8.1f p999=%8.1f max=%9.1f" % (f.split("/")[-1], r["rps"], lat["p50"], lat["p90"], lat["p99"], lat["p999"], lat["max"]))
EOF
for f in "$OUT"/profile_*.txt; do echo "== $f"; cut -c1-120 "$f"; done
