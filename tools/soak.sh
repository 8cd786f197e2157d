#!/usr/bin/env bash
# Mixed-traffic soak: /api/hola at full tilt, /api/hola-chaining against a dead downstream (breaker
# cycling open/half-open), short-lived /hystrix.stream subscribers, config file in use, tracing to a
# dead collector - while watching the server's resident memory and descriptor count.
#   tools/soak.sh [seconds]        (default 60)
set -u
cd "$(dirname "$0")/.."
SECS=${1:-60}
PORT=18090
env -u HOSTNAME ZIPKIN_SERVER_URL=http://127.0.0.1:9 ./build/hola --port $PORT --bind 127.0.0.1 --threads 4 \
    --aloha-url http://127.0.0.1:9/ --quiet -Dconf=tests/data/translation.properties &
SRV=$!
sleep 0.3
rss() { awk '/VmRSS/ {print $2}' /proc/$SRV/status; }
fds() { ls /proc/$SRV/fd | wc -l; }
echo "t=0s rss=$(rss)kB fds=$(fds)"
./build/hola-load --json -c 64 -d "$SECS" -w 0 -t 2 http://127.0.0.1:$PORT/api/hola > /tmp/soak_hola.json &
L1=$!
./build/hola-load --json -c 32 -d "$SECS" -w 0 -t 1 http://127.0.0.1:$PORT/api/hola-chaining > /tmp/soak_chain.json &
L2=$!
( end=$((SECONDS + SECS)); while [ $SECONDS -lt $end ]; do curl -s -m 0.2 "http://127.0.0.1:$PORT/hystrix.stream?delay=20" > /dev/null; done ) &
L3=$!
for frac in 1 2 3 4; do
  sleep $((SECS / 5))
  echo "t=$((frac * SECS / 5))s rss=$(rss)kB fds=$(fds)"
done
wait $L1 $L2 $L3 2>/dev/null
sleep 1
echo "t=end rss=$(rss)kB fds=$(fds)  (load finished, connections closed)"
python - <<'EOF'
import json
for name in ("hola", "chain"):
    r = json.load(open("/tmp/soak_%s.json" % name))
    print("%-6s requests=%d rps=%.0f p50=%.1fus p99=%.1fus errors=%d" % (name, r["requests"], r["rps"], r["latency_us"]["p50"], r["latency_us"]["p99"], r["errors"]["total"]))
EOF
kill $SRV
wait $SRV
echo "server exit code: $?"
