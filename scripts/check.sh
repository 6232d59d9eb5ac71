#!/usr/bin/env bash
# Everything that can be checked without a GPU, in the order a change should pass it:
#   scripts/check.sh              build (sm_100a cross-compile), native unit suites, CPU pytest
#   scripts/check.sh --sanitize   additionally: AddressSanitizer and ThreadSanitizer builds of the daemons and libocm,
#                                 native clients (ocm_test, ocm_mt, hello_ocm) against sanitized daemons
# GPU side: python -m pytest tests -q -m gpu ; python bench.py   
set -euo pipefail
cd "$(dirname "$0")/.."
ROOT=$PWD

echo "== build"
python -c "import __graft_entry__ as g; g.build()"
echo "== native unit suites"
build/bin/unit_host all | tail -n 12
echo "== pytest (CPU)"
python -m pytest tests -x -q -m "not gpu"

if [[ "${1:-}" != "--sanitize" ]]; then exit 0; fi

CXX=${CXX_SAN:-/usr/bin/g++}
CC=${CC_SAN:-/usr/bin/gcc}
for san in address thread; do
    echo "== ${san} sanitizer"
    B=$(mktemp -d /tmp/ocm_${san}_XXXX)
    cmake -S "$ROOT" -B "$B" -G Ninja -DOCM_SANITIZE=$san -DCMAKE_CXX_COMPILER=$CXX -DCMAKE_C_COMPILER=$CC > "$B/configure.log"
    cmake --build "$B" --target oncillamem ocm ocm_test ocm_mt hello_ocm unit_host > "$B/build.log"
    "$B/bin/unit_host" all > "$B/unit.log" 2>&1
    NF="$B/nodefile"
    printf '0 localhost 127.0.0.1 %d 0 host_bytes=32M\n1 localhost 127.0.0.1 %d 0 host_bytes=32M\n' $((20000 + RANDOM % 5000)) $((25000 + RANDOM % 4000)) > "$NF"
    export OCM_NS="san$$" OCM_NO_GPU=1 OCM_LOG=warn ASAN_OPTIONS=detect_leaks=0 TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0"
    OCM_RANK=0 "$B/bin/oncillamem" "$NF" > "$B/d0.log" 2>&1 & D0=$!
    sleep 2
    OCM_RANK=1 "$B/bin/oncillamem" "$NF" > "$B/d1.log" 2>&1 & D1=$!
    sleep 2
    rc=0
    for args in "1 1 2 1" "2 1 1 3" "3 1 1 1" "5 1 1 8"; do
        OCM_RANK=0 timeout 120 "$B/bin/ocm_test" $args >> "$B/clients.log" 2>&1 || rc=1
        OCM_RANK=0 OCM_ALLOC_PATH=daemon timeout 120 "$B/bin/ocm_test" $args >> "$B/clients.log" 2>&1 || rc=1
    done
    OCM_RANK=1 timeout 300 "$B/bin/ocm_mt" 6 40 >> "$B/clients.log" 2>&1 || rc=1
    OCM_RANK=1 timeout 60 "$B/bin/hello_ocm" >> "$B/clients.log" 2>&1 || rc=1
    kill -TERM $D1 $D0
    wait $D1 $D0 2>/dev/null || true
    reports=$(cat "$B"/*.log | grep -c -E "ERROR: AddressSanitizer|WARNING: ThreadSanitizer" || true)
    echo "   clients rc=$rc, sanitizer reports=$reports (logs in $B)"
    if [[ $rc -ne 0 || $reports -ne 0 ]]; then exit 1; fi
    rm -rf "$B"
done
echo "== all checks passed"
