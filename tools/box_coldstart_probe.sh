#!/usr/bin/env bash
# Why does exec -> first 200 sometimes take 20-40 ms on the box? 40 launches per variant.
set -u
OUT=gpurun_out/r2_coldstart
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
make static > /dev/null 2>&1
git_old=/tmp/hola_old
rm -rf $git_old; mkdir -p $git_old
run() { label=$1; shift; env -u HOSTNAME -u ZIPKIN_SERVER_URL ./build/hola-coldstart --port 18084 --runs 40 -- "$@" > "$OUT/$label.json" 2>&1
  python - "$label" "$OUT/$label.json" <<'PY'
import json,sys
try:
    r=json.load(open(sys.argv[2])); xs=sorted(x/1000 for x in r["runs"])
    print("%-28s min %.2f median %.2f p90 %.2f max %.2f  over5ms %d/%d" % (sys.argv[1], xs[0], xs[len(xs)//2], xs[int(len(xs)*0.9)], xs[-1], sum(x>5 for x in xs), len(xs)))
except Exception as e:
    print(sys.argv[1], "failed", e, open(sys.argv[2]).read()[:200])
PY
}
B="--port 18084 --bind 127.0.0.1 --quiet"
run dynamic_auto        ./build/hola $B
run dynamic_epoll       ./build/hola $B --io-backend epoll
run static_auto         ./build/hola-static $B
run static_epoll        ./build/hola-static $B --io-backend epoll
run dynamic_auto_t4     ./build/hola $B --threads 4
run dynamic_epoll_t4    ./build/hola $B --threads 4 --io-backend epoll
run dynamic_epoll_reuse ./build/hola $B --io-backend epoll --accept-mode reuseport
run dynamic_epoll_rr    ./build/hola $B --io-backend epoll --accept-mode rr
run dynamic_auto_again  ./build/hola $B
run static_auto_again   ./build/hola-static $B
