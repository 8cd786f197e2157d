set -u
OUT=gpurun_out/r2_coldstart32; mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
run() { label=$1; shift; env -u HOSTNAME -u ZIPKIN_SERVER_URL ./build/hola-coldstart --port 18084 --runs 20 -- "$@" > "$OUT/$label.json" 2>&1
  python - "$label" "$OUT/$label.json" <<'PY'
import json,sys
try:
    r=json.load(open(sys.argv[2])); xs=sorted(x/1000 for x in r["runs"])
    print("%-22s min %.2f median %.2f p90 %.2f max %.2f  over5ms %d/%d" % (sys.argv[1], xs[0], xs[len(xs)//2], xs[int(len(xs)*0.9)], xs[-1], sum(x>5 for x in xs), len(xs)))
except Exception as e:
    print(sys.argv[1], "failed", e)
PY
}
B="--port 18084 --bind 127.0.0.1 --quiet"
nproc
run auto_default   ./build/hola $B
run epoll_default  ./build/hola $B --io-backend epoll
run auto_t16       ./build/hola $B --threads 16
run epoll_rr       ./build/hola $B --io-backend epoll --accept-mode rr
run epoll_reuse    ./build/hola $B --io-backend epoll --accept-mode reuseport
