#!/usr/bin/env bash
# Mixed-traffic soak: /api/hola at full tilt, /api/hola-chaining against a dead downstream (breaker
# cycling open/half-open), short-lived /hystrix.stream subscribers, config file in use, tracing to a
# dead collector, connection churn (clients sending "Connection: close": a new connection per request, so closed
# connections with io_uring operations in flight come and go all the time) and a second instance
# whose downstream is alive (the outbound client's pooled connections) - while watching the
# servers' resident memory and descriptor counts.
#   tools/soak.sh [seconds]        (default 60)
set -u
cd "$(dirname "$0")/.."
SECS=${1:-60}
PORT=18090
env -u HOSTNAME ZIPKIN_SERVER_URL=http://127.0.0.1:9 ./build/hola --port $PORT --bind 127.0.0.1 --threads 4 \
    --aloha-url http://127.0.0.1:9/ --quiet -Dconf=tests/data/translation.properties &
SRV=$!
env -u HOSTNAME -u ZIPKIN_SERVER_URL ./build/hola --port $((PORT + 1)) --bind 127.0.0.1 --threads 2 --quiet --serve-aloha &
ALOHA=$!
env -u HOSTNAME -u ZIPKIN_SERVER_URL ./build/hola --port $((PORT + 2)) --bind 127.0.0.1 --threads 2 --quiet \
    --aloha-url http://127.0.0.1:$((PORT + 1))/ -Dhystrix.threadpool.default.coreSize=64 &
SRV2=$!
sleep 0.3
rss() { echo "$(awk '/VmRSS/ {print $2}' /proc/$SRV/status)/$(awk '/VmRSS/ {print $2}' /proc/$SRV2/status)"; }
fds() { echo "$(ls /proc/$SRV/fd | wc -l)/$(ls /proc/$SRV2/fd | wc -l)"; }
echo "t=0s rss=$(rss)kB fds=$(fds)   (first/second instance)"
./build/hola-load --json -c 64 -d "$SECS" -w 0 -t 2 http://127.0.0.1:$PORT/api/hola > /tmp/soak_hola.json &
L1=$!
./build/hola-load --json -c 32 -d "$SECS" -w 0 -t 1 http://127.0.0.1:$PORT/api/hola-chaining > /tmp/soak_chain.json &
L2=$!
./build/hola-load --json -c 8 -d "$SECS" -w 0 -t 1 -H "Connection: close" http://127.0.0.1:$PORT/api/hola > /tmp/soak_churn.json &
L4=$!
./build/hola-load --json -c 32 -d "$SECS" -w 0 -t 1 http://127.0.0.1:$((PORT + 2))/api/hola-chaining > /tmp/soak_live.json &
L5=$!
( end=$((SECONDS + SECS)); while [ $SECONDS -lt $end ]; do curl -s -m 0.2 "http://127.0.0.1:$PORT/hystrix.stream?delay=20" > /dev/null; done ) &
L3=$!
for frac in 1 2 3 4; do
  sleep $((SECS / 5))
  echo "t=$((frac * SECS / 5))s rss=$(rss)kB fds=$(fds)"
done
wait $L1 $L2 $L3 $L4 $L5 2>/dev/null
sleep 1
echo "t=end rss=$(rss)kB fds=$(fds)  (load finished, connections closed)"
python - <<'EOF'
import json
for name in ("hola", "chain", "churn", "live"):
    r = json.load(open("/tmp/soak_%s.json" % name))
    print("%-6s requests=%d rps=%.0f p50=%.1fus p99=%.1fus errors=%d" % (name, r["requests"], r["rps"], r["latency_us"]["p50"], r["latency_us"]["p99"], r["errors"]["total"]))
EOF
kill -USR1 $SRV; sleep 0.2
kill $SRV $SRV2 $ALOHA
wait $SRV; echo "server exit code: $?"
wait $SRV2; echo "second instance exit code: $?"
