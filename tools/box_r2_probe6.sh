#!/usr/bin/env bash
# Round 2, sixth pass (1-GPU box): every outcome of /api/hola-chaining again, now that the outbound
# client rides on io_uring too, on both backends; plus a bench sanity run. Results in gpurun_out/r2_probe6/.
set -u
OUT=gpurun_out/r2_probe6
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build()" > "$OUT/build.txt" 2>&1
python tools/chaining_paths.py --seconds 8 --out "$OUT/chaining_uring.json" > "$OUT/chaining_uring.txt" 2>&1
cat "$OUT/chaining_uring.txt"
python bench.py --steps 20 --warmup 3 > "$OUT/bench_n1.json" 2> "$OUT/bench.err"
python -c "
import json; r=json.loads(open('$OUT/bench_n1.json').read().strip().splitlines()[-1])
print('bench', round(r['value']), r['per_server_core_rps'], r['latency_us'], r['config']['io_backend'], r['coldstart_ms'], r['half_load'])"
python -m pytest tests/test_chaining.py tests/test_tracing.py tests/test_stream.py -q -x > "$OUT/pytest.txt" 2>&1; tail -2 "$OUT/pytest.txt"
