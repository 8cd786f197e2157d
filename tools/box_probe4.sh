#!/usr/bin/env bash
# Fourth pass on the benchmark box, after the request-parser rework: headline again, and the same
# load with browser-sized request heads (11 header lines, ~560 bytes) to see what realistic
# clients cost. Results in gpurun_out/probe4/.
set -u
OUT=gpurun_out/probe4
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build()" > "$OUT/build.txt" 2>&1
HDRS=(-H "User-Agent: Mozilla/5.0 (X11; Linux x86_64) AppleWebKit/537.36 (KHTML, like Gecko) Chrome/120.0 Safari/537.36"
      -H "Accept: text/html,application/xhtml+xml,application/xml;q=0.9,image/webp,*/*;q=0.8"
      -H "Accept-Language: en-US,en;q=0.5" -H "Accept-Encoding: gzip, deflate" -H "Connection: keep-alive"
      -H "Cookie: session=abcdef0123456789abcdef0123456789; theme=dark" -H "Cache-Control: no-cache" -H "Pragma: no-cache"
      -H "X-Requested-With: XMLHttpRequest" -H "Referer: http://localhost:8080/")
run() {  # run <label> <hola-load args...>
  local label=$1; shift
  env -u HOSTNAME ZIPKIN_SERVER_URL=http://127.0.0.1:9 ./build/hola --port 18080 --bind 127.0.0.1 --threads 8 --cpus 0-7 --verbose 2> "$OUT/server_$label.err" &
  local pid=$!
  sleep 0.3
  ./build/hola-load --json -c 64 -d 10 -w 2 -t 8 --cpus 8-15 --expect-body "Hola de unknown" "$@" http://127.0.0.1:18080/api/hola > "$OUT/$label.json" 2>> "$OUT/server_$label.err"
  kill $pid; wait $pid 2>/dev/null
  grep "stats:" "$OUT/server_$label.err" | sed 's/.*stats: //' > "$OUT/profile_$label.txt"
}
run minimal_head
run browser_head "${HDRS[@]}"
run minimal_head_pipe16 --pipeline 16
run browser_head_pipe16 --pipeline 16 "${HDRS[@]}"
python bench.py --steps 30 --warmup 3 > "$OUT/bench_n1_30s.json" 2> "$OUT/bench.err"
python - <<'EOF'
import json, glob
for f in sorted(glob.glob("gpurun_out/probe4/*.json")):
    r = json.loads(open(f).read().strip().splitlines()[-1])
    lat = r["latency_us"]
    print("%-28s rps=%9d p50=%7.1f p99=%7.1f" % (f.split("/")[-1], r.get("rps", r.get("value")), lat["p50"], lat["p99"]))
EOF
for f in "$OUT"/profile_*.txt; do echo "== $f"; cut -c1-130 "$f"; done
