#!/usr/bin/env bash
# Quick end-of-change verification on the benchmark box: what the driver will do at round end
# (build, smoke, pytest -m gpu, bench reference arm, bench N=1), plus the full CPU test-suite.
# Results in gpurun_out/verify/.
set -u
OUT=gpurun_out/verify
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > "$OUT/build_smoke.txt" 2>&1; echo "rc=$?" >> "$OUT/build_smoke.txt"
python -m pytest tests -x -q -m gpu -p no:cacheprovider > "$OUT/pytest_gpu.txt" 2>&1; echo "rc=$?" >> "$OUT/pytest_gpu.txt"
python -m pytest tests -x -q -m "not gpu" -p no:cacheprovider > "$OUT/pytest_not_gpu.txt" 2>&1; echo "rc=$?" >> "$OUT/pytest_not_gpu.txt"
python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > "$OUT/bench_reference.json" 2>&1; echo "rc=$?" >> "$OUT/bench_reference.json"
python bench.py --gpus 1 --steps 20 --warmup 3 > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"; echo "rc=$?" >> "$OUT/bench_n1.err"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29714 bench.py --gpus 4 --steps 5 --warmup 3 > "$OUT/bench_n4.json" 2> "$OUT/bench_n4.err"; echo "rc=$?" >> "$OUT/bench_n4.err"
for f in build_smoke.txt pytest_gpu.txt pytest_not_gpu.txt bench_reference.json; do echo "== $f"; tail -n 3 "$OUT/$f"; done
python - <<'EOF'
import json
for f in ("bench_n1", "bench_n4"):
    r = json.loads(open("gpurun_out/verify/%s.json" % f).read().strip().splitlines()[-1])
    print(f, r["n_gpus"], r["value"], r["latency_us"], r["gpu_launches"], r["clocks"])
EOF
