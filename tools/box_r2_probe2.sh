#!/usr/bin/env bash
# Round 2, second pass on the benchmark box: the measured ceiling (hola vs a null server, epoll vs
# io_uring), tracing with a live collector, and every outcome of /api/hola-chaining.
# Results in gpurun_out/r2_probe2/.
set -u
OUT=gpurun_out/r2_probe2
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build()" > "$OUT/build.txt" 2>&1; echo "build rc=$?" >> "$OUT/build.txt"
cat /proc/sys/kernel/io_uring_disabled > "$OUT/io_uring_disabled.txt" 2>&1
python tools/ceiling.py --seconds 8 --repeat 3 --out "$OUT/ceiling.json" > "$OUT/ceiling.txt" 2>&1
cat "$OUT/ceiling.txt"
python tools/chaining_paths.py --seconds 8 --out "$OUT/chaining.json" > "$OUT/chaining.txt" 2>&1
cat "$OUT/chaining.txt"
python tools/tracing_cost.py --seconds 5 --out "$OUT/tracing.json" > "$OUT/tracing.txt" 2>&1
cat "$OUT/tracing.txt"
