#!/usr/bin/env bash
# Last check of HEAD on the benchmark box: build + smoke, unit tests, the whole pytest suite, the
# `-m gpu` selection, one bench run (20 s) and the reference arm. Results in gpurun_out/r2_verify/.
set -u
OUT=gpurun_out/r2_verify
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > "$OUT/build_smoke.txt" 2>&1; echo "build+smoke rc=$?" | tee -a "$OUT/build_smoke.txt"
build/hola_tests > "$OUT/unit.txt" 2>&1; tail -1 "$OUT/unit.txt"
python -m pytest tests -x -q -m "not gpu" > "$OUT/pytest_not_gpu.txt" 2>&1; tail -2 "$OUT/pytest_not_gpu.txt"
python -m pytest tests -x -q -m gpu > "$OUT/pytest_gpu.txt" 2>&1; echo "gpu selection rc=$?"; tail -2 "$OUT/pytest_gpu.txt"
python bench.py --steps 20 --warmup 3 > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"
python bench.py --impl reference --steps 20 --warmup 3 > "$OUT/bench_reference.json" 2>&1
python -c "
import json; r=json.loads(open('$OUT/bench_n1.json').read().strip().splitlines()[-1])
print('bench', round(r['value']), r['per_server_core_rps'], r['latency_us'], r['config']['io_backend'], r['coldstart_ms'], r['half_load'], r['worker_spread'])"
cat "$OUT/bench_reference.json"
