#!/bin/bash
# The CPU test-suite against an AddressSanitizer + UBSan build of the host runtime: the Python tests reach
# paths the C++ sanitizer driver does not (batch packers, page-view merge, plugins, the dataset API as users
# call it).  The instrumented library is built out of tree and loaded through PBX_HOST_LIB; libasan is
# preloaded because the interpreter itself is not instrumented.
#   benchmarks/sanitize_pytests.sh [outdir] [pytest args...]      -> <outdir>/sanitizer_pytests_asan.log
set -u
cd "$(dirname "$0")/.."
OUT=${1:-profiles}
mkdir -p "$OUT"
shift || true
ARGS=${*:-tests/test_host_runtime.py tests/test_api_sweep.py tests/test_cpu_e2e.py tests/test_graph.py tests/test_legacy_feed.py tests/test_pass_build.py tests/test_properties.py}
TMP=$(mktemp -d)
g++ -std=c++17 -O1 -g -fno-omit-frame-pointer -fsanitize=address,undefined -fno-sanitize-recover=undefined \
    -fPIC -shared -pthread -Iinclude -o $TMP/libpbx_host_asan.so csrc/host/*.cc -ldl || { echo "build failed"; exit 1; }
LIBASAN=$(g++ -print-file-name=libasan.so)
LIBUBSAN=$(g++ -print-file-name=libubsan.so)
{
  echo "# pytest $ARGS with libpbx_host built -fsanitize=address,undefined"
  # (leak detection off: the interpreter and torch keep memory until exit; python's own allocator is left alone)
  PBX_HOST_LIB=$TMP/libpbx_host_asan.so LD_PRELOAD="$LIBASAN $LIBUBSAN" \
  ASAN_OPTIONS=detect_leaks=0:abort_on_error=1:halt_on_error=1 UBSAN_OPTIONS=print_stacktrace=1 \
    python -m pytest $ARGS -x -q -m "not gpu" -p no:cacheprovider
  echo "exit code $?"
} > $OUT/sanitizer_pytests_asan.log 2>&1
tail -4 $OUT/sanitizer_pytests_asan.log
rc=1; grep -q "exit code 0" $OUT/sanitizer_pytests_asan.log && rc=0
rm -rf $TMP
exit $rc
