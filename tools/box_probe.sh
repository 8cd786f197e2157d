#!/usr/bin/env bash
# Run on the benchmark box (via gpurun): records what the machine is, checks that the project
# builds and passes its tests there, and sweeps the server/generator CPU split so that bench.py's
# defaults can be chosen from measurements. Everything lands in gpurun_out/.
set -u
OUT=gpurun_out
mkdir -p "$OUT"
{
  echo "== date"; date -u
  echo "== cpu"; nproc; lscpu | grep -E "Model name|Socket|Core|Thread|NUMA|MHz|L2|L3"
  echo "== mem"; free -g | head -2
  echo "== kernel"; uname -r
  echo "== gpu"; nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader 2>&1 | head -8
  echo "== java/mvn/docker (reference arm prerequisites)"; for c in java javac mvn docker podman; do printf "%s: " $c; command -v $c || echo absent; done
} > "$OUT/box.txt" 2>&1

python -c "import __graft_entry__ as g; g.build(); g.smoke()" > "$OUT/smoke.txt" 2>&1; echo "smoke rc=$?" >> "$OUT/smoke.txt"
./build/hola_tests > "$OUT/cpp_tests.txt" 2>&1; echo "rc=$?" >> "$OUT/cpp_tests.txt"
python -m pytest tests -q -x -m "not gpu" -p no:cacheprovider > "$OUT/pytest_not_gpu.txt" 2>&1; echo "rc=$?" >> "$OUT/pytest_not_gpu.txt"
python -m pytest tests -q -x -m gpu -p no:cacheprovider > "$OUT/pytest_gpu.txt" 2>&1; echo "rc=$?" >> "$OUT/pytest_gpu.txt"

# Sweep: S server threads on CPUs [0,S), L generator threads on CPUs [S,S+L).
NCPU=$(nproc)
sweep() {
  local S=$1 L=$2 MODE=$3
  local last=$((S+L-1))
  if [ "$last" -ge "$NCPU" ]; then return; fi
  local envz=()
  if [ "$MODE" = "trace" ]; then envz=(ZIPKIN_SERVER_URL=http://127.0.0.1:9); fi
  env -u HOSTNAME -u ZIPKIN_SERVER_URL "${envz[@]}" ./build/hola --port 18080 --bind 127.0.0.1 --threads "$S" --cpus "0-$((S-1))" --quiet &
  local pid=$!
  sleep 0.3
  printf '{"server_threads":%d,"load_threads":%d,"mode":"%s","report":' "$S" "$L" "$MODE"
  ./build/hola-load --json -c 64 -d 4 -w 1 -t "$L" --cpus "$S-$last" --expect-body "Hola de unknown" http://127.0.0.1:18080/api/hola | python -c "import json,sys; r=json.load(sys.stdin); print(json.dumps({k:r[k] for k in ('rps','latency_us','errors')}), end='')"
  echo '}'
  kill $pid; wait $pid 2>/dev/null
}
: > "$OUT/sweep.jsonl"
for cfg in "1 1" "2 2" "4 4" "8 8" "16 16" "32 32" "4 8" "8 16" "16 32" "8 4" "16 8" "32 16" "24 24" "12 12"; do
  set -- $cfg
  sweep "$1" "$2" notrace >> "$OUT/sweep.jsonl" 2>> "$OUT/sweep.err"
done
for cfg in "4 4" "8 8" "16 16"; do
  set -- $cfg
  sweep "$1" "$2" trace >> "$OUT/sweep.jsonl" 2>> "$OUT/sweep.err"
done

python bench.py --steps 10 --warmup 3 > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"; echo "rc=$?" >> "$OUT/bench_n1.err"
python bench.py --steps 10 --warmup 3 --mode notrace > "$OUT/bench_n1_notrace.json" 2>> "$OUT/bench_n1.err"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > "$OUT/bench_n2.json" 2> "$OUT/bench_n2.err"; echo "rc=$?" >> "$OUT/bench_n2.err"
python bench.py --impl reference --gpus 1 --steps 1 --warmup 1 > "$OUT/bench_reference.json" 2>&1
tail -3 "$OUT"/pytest_*.txt "$OUT/smoke.txt" "$OUT/cpp_tests.txt"
cat "$OUT/bench_n1.json"
