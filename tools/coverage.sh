#!/usr/bin/env bash
# Line coverage of the native code under the whole test-suite (C++ unit tests + the black-box
# pytest files run against the instrumented executable). Prints a per-file table and writes
# profiles/coverage.txt. Needs gcov (part of the distribution compiler).
set -euo pipefail
cd "$(dirname "$0")/.."
make cov -j"$(nproc)" > /dev/null
find build/cov -name '*.gcda' -delete
./build/cov/hola_tests > /dev/null
python -m pytest tests/test_contract.py tests/test_server_core.py tests/test_chaining.py tests/test_tracing.py \
    tests/test_stream.py tests/test_dns.py tests/test_fuzz.py tests/test_cli.py -q --variant cov -p no:cacheprovider > /dev/null
# The building blocks through the C ABI (hypothesis tests), against the instrumented library.
HOLA_LIB_VARIANT=cov python -m pytest tests/test_native_units.py -q -p no:cacheprovider > /dev/null
out=profiles/coverage.txt
{
  echo "# gcov line coverage, $(date -u +%Y-%m-%d), C++ unit tests + black-box pytest against build/cov/hola + C-ABI tests against build/cov/libhola.so"
  printf '%-32s %8s %8s\n' file lines covered
  total=0; hit=0
  for src in src/core/*.cc src/net/*.cc src/app/*.cc; do
    obj="build/cov/obj/${src#src/}"; obj="${obj%.cc}.o"
    [ -f "${obj%.o}.gcda" ] || { printf '%-32s %8s %8s\n' "$src" "-" "not run"; continue; }
    line=$(gcov -n -o "$(dirname "$obj")" "$src" 2>/dev/null | grep -A1 "File '$src'" | tail -1)
    pct=$(echo "$line" | sed -E 's/Lines executed:([0-9.]+)% of ([0-9]+)/\1 \2/')
    p=${pct% *}; n=${pct#* }
    printf '%-32s %8s %7s%%\n' "$src" "$n" "$p"
    total=$((total + n)); hit=$(python -c "print($hit + round($n * $p / 100))")
  done
  python -c "print('%-32s %8d %7.1f%%' % ('TOTAL', $total, 100.0 * $hit / max(1, $total)))"
} | tee "$out"
