#!/usr/bin/env bash
# Third measurement pass on the benchmark box: latency below saturation (open loop, no
# coordinated omission), the server's headroom when the kernel cost is amortised (pipelining),
# and the full test-suite including the sanitizer variants. Results in gpurun_out/probe3/.
set -u
OUT=gpurun_out/probe3
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build()" > "$OUT/build.txt" 2>&1

serve() {  # serve <label> <extra env...> ; starts the server on 8 cores, traced like the headline run
  env -u HOSTNAME ZIPKIN_SERVER_URL=http://127.0.0.1:9 ./build/hola --port 18080 --bind 127.0.0.1 --threads 8 --cpus 0-7 --verbose 2> "$OUT/server_$1.err" &
  SRV=$!
  sleep 0.3
}
stop() { kill $SRV; wait $SRV 2>/dev/null; grep "stats:" "$OUT/server_$1.err" | sed 's/.*stats: //' > "$OUT/profile_$1.txt"; }

# 1. Open loop at fixed rates: what a client sees when the service is not saturated.
for RATE in 250000 500000 1000000 1500000 2000000; do
  serve open_$RATE
  ./build/hola-load --json -c 64 -d 8 -w 2 -t 8 --cpus 8-15 --rate $RATE --expect-body "Hola de unknown" http://127.0.0.1:18080/api/hola > "$OUT/open_$RATE.json" 2>> "$OUT/server_open_$RATE.err"
  stop open_$RATE
done

# 2. Pipelined clients (not the headline configuration): server capacity with the per-request
#    kernel cost amortised over 16 requests per read/write.
for P in 4 16; do
  serve pipe_$P
  ./build/hola-load --json -c 64 -d 8 -w 2 -t 8 --cpus 8-15 --pipeline $P --expect-body "Hola de unknown" http://127.0.0.1:18080/api/hola > "$OUT/pipeline_$P.json" 2>> "$OUT/server_pipe_$P.err"
  stop pipe_$P
done

# 3. Many connections.
serve conns_4096
./build/hola-load --json -c 4096 -d 8 -w 2 -t 8 --cpus 8-15 --expect-body "Hola de unknown" http://127.0.0.1:18080/api/hola > "$OUT/conns_4096.json" 2>> "$OUT/server_conns_4096.err"
stop conns_4096

# 4. Whole test-suite, then the black-box part again under the sanitizers.
python -m pytest tests -q -p no:cacheprovider > "$OUT/pytest_all.txt" 2>&1; echo "rc=$?" >> "$OUT/pytest_all.txt"
python -m pytest tests/test_contract.py tests/test_server_core.py tests/test_chaining.py tests/test_tracing.py tests/test_stream.py tests/test_dns.py tests/test_fuzz.py -q -p no:cacheprovider --variant asan > "$OUT/pytest_asan.txt" 2>&1; echo "rc=$?" >> "$OUT/pytest_asan.txt"
python -m pytest tests/test_contract.py tests/test_server_core.py tests/test_chaining.py tests/test_tracing.py tests/test_stream.py tests/test_dns.py tests/test_fuzz.py -q -p no:cacheprovider --variant tsan > "$OUT/pytest_tsan.txt" 2>&1; echo "rc=$?" >> "$OUT/pytest_tsan.txt"

python - <<'EOF'
import json, glob
for f in sorted(glob.glob("gpurun_out/probe3/*.json")):
    try:
        r = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    lat = r["latency_us"]
    print("%-22s rps=%9d p50=%8.1f p99=%8.1f p999=%8.1f max=%9.1f errors=%d" % (f.split("/")[-1], r["rps"], lat["p50"], lat["p99"], lat["p999"], lat["max"], r["errors"]["total"]))
EOF
for f in "$OUT"/profile_pipe_*.txt "$OUT"/profile_conns_4096.txt; do echo "== $f"; cat "$f"; done
for f in "$OUT"/pytest_*.txt; do echo "== $f"; tail -n 3 "$f"; done
