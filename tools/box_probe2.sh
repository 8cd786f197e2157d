#!/usr/bin/env bash
# Second measurement pass on the benchmark box: the full 60 s headline run, the 1/2/4/8-instance
# series the driver will run, CPU profile of the server under load, accept-mode comparison,
# cold start, syscall cost. Results in gpurun_out/probe2/.
set -u
OUT=gpurun_out/probe2
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build()" > "$OUT/build.txt" 2>&1

# 1. Raw syscall cost on this machine (context for "how much of a request is kernel time").
cat > /tmp/sysc.c <<'EOF'
#include <stdio.h>
#include <unistd.h>
#include <sys/syscall.h>
#include <time.h>
static double now(){struct timespec t; clock_gettime(CLOCK_MONOTONIC,&t); return t.tv_sec*1e9+t.tv_nsec;}
int main(){ int N=2000000; double t0=now(); for(int i=0;i<N;i++) syscall(SYS_getppid); double t1=now(); printf("getppid: %.0f ns/call\n",(t1-t0)/N);
 t0=now(); for(int i=0;i<N;i++){struct timespec t; clock_gettime(CLOCK_MONOTONIC,&t);} t1=now(); printf("clock_gettime: %.0f ns/call\n",(t1-t0)/N); return 0;}
EOF
gcc -O2 -o /tmp/sysc /tmp/sysc.c && /tmp/sysc > "$OUT/syscall_cost.txt" 2>&1

# 2. Server CPU profile under the headline load (8 server threads, 8 generator threads).
profile() {
  local label=$1; shift
  env -u HOSTNAME -u ZIPKIN_SERVER_URL "$@" ./build/hola --port 18080 --bind 127.0.0.1 --threads 8 --cpus 0-7 --verbose $EXTRA 2> "$OUT/server_$label.err" &
  local pid=$!
  sleep 0.3
  ./build/hola-load --json -c 64 -d 8 -w 2 -t 8 --cpus 8-15 --expect-body "Hola de unknown" http://127.0.0.1:18080/api/hola > "$OUT/load_$label.json" 2>> "$OUT/server_$label.err"
  kill $pid; wait $pid 2>/dev/null
  grep "stats:" "$OUT/server_$label.err" | sed 's/.*stats: //' > "$OUT/profile_$label.txt"
}
EXTRA="" profile notrace
EXTRA="" profile trace ZIPKIN_SERVER_URL=http://127.0.0.1:9
EXTRA="--accept-mode rr" profile notrace_rr
EXTRA="--spin-us 50" profile notrace_spin50

# 3. Cold start.
env -u HOSTNAME -u ZIPKIN_SERVER_URL ./build/hola-coldstart --port 18084 --runs 15 -- ./build/hola --port 18084 --bind 127.0.0.1 --threads 8 > "$OUT/coldstart_dynamic.json" 2>&1
make static > /dev/null 2>&1
env -u HOSTNAME -u ZIPKIN_SERVER_URL ./build/hola-coldstart --port 18084 --runs 15 -- ./build/hola-static --port 18084 --bind 127.0.0.1 --threads 8 > "$OUT/coldstart_static.json" 2>&1

# 4. The headline run exactly as BASELINE.json config 4 words it: 64 clients, 60 s.
python bench.py --steps 60 --warmup 3 > "$OUT/bench_n1_60s.json" 2> "$OUT/bench_n1_60s.err"
python bench.py --steps 20 --warmup 3 --mode conf > "$OUT/bench_n1_conf.json" 2>> "$OUT/bench_n1_60s.err"

# 5. The scaling series the driver runs (torchrun for N > 1).
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 10 --warmup 3 > "$OUT/bench_n$N.json" 2> "$OUT/bench_n$N.err"
done

# 6. Chaining endpoint throughput (fallback path: downstream unresolvable, circuit open most of the time).
env -u HOSTNAME -u ZIPKIN_SERVER_URL ./build/hola --port 18080 --bind 127.0.0.1 --threads 8 --cpus 0-7 --quiet --aloha-url http://127.0.0.1:9/ &
pid=$!; sleep 0.3
./build/hola-load --json -c 64 -d 5 -w 1 -t 8 --cpus 8-15 --expect-body '["Hola de unknown","Aloha response (fallback)"]' http://127.0.0.1:18080/api/hola-chaining > "$OUT/load_chaining_fallback.json" 2>&1
kill $pid; wait $pid 2>/dev/null

for f in "$OUT"/profile_*.txt "$OUT"/coldstart_*.json "$OUT/syscall_cost.txt"; do echo "== $f"; cat "$f"; done
python - <<'EOF'
import json, glob
for f in sorted(glob.glob("gpurun_out/probe2/bench_n*.json")) + sorted(glob.glob("gpurun_out/probe2/load_*.json")):
    try:
        r = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    lat = r.get("latency_us", {})
    print(f.split("/")[-1], r.get("n_gpus", ""), round(r.get("value", r.get("rps", 0))), lat.get("p50"), lat.get("p99"), r.get("errors"))
EOF
