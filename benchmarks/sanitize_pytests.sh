#!/bin/bash
# The CPU test-suite against an AddressSanitizer + UBSan build of the host runtime: the Python tests reach
# paths the C++ sanitizer driver does not (batch packers, page-view merge, plugins, the dataset API as users
# call it).  The instrumented library is built out of tree and loaded through PBX_HOST_LIB; libasan is
# preloaded because the interpreter itself is not instrumented.
#   benchmarks/sanitize_pytests.sh [outdir] [pytest args...]      -> <outdir>/sanitizer_pytests_asan.log
#   SAN=tsan benchmarks/sanitize_pytests.sh ...                   -> <outdir>/sanitizer_pytests_tsan.log
# (tsan: tests that read through a pipe_command are left out -- the child shell would inherit the preload -- and so is
#  the 5 MB-line reader test, whose shadow memory does not fit this container)
set -u
cd "$(dirname "$0")/.."
OUT=${1:-profiles}
mkdir -p "$OUT"
shift || true
ARGS=${*:-tests/test_host_runtime.py tests/test_api_sweep.py tests/test_cpu_e2e.py tests/test_graph.py tests/test_legacy_feed.py tests/test_pass_build.py tests/test_properties.py tests/test_index_dataset.py}
TMP=$(mktemp -d)
SAN=${SAN:-asan}
if [ $SAN = tsan ]; then
  FL="-fsanitize=thread"; PRE=$(g++ -print-file-name=libtsan.so); KEXPR="not pipe_command and not block_boundaries"
else
  FL="-fsanitize=address,undefined -fno-sanitize-recover=undefined"
  PRE="$(g++ -print-file-name=libasan.so) $(g++ -print-file-name=libubsan.so)"; KEXPR=""
fi
g++ -std=c++17 -O1 -g -fno-omit-frame-pointer $FL -fPIC -shared -pthread -Iinclude \
    -o $TMP/libpbx_host_$SAN.so csrc/host/*.cc -ldl || { echo "build failed"; exit 1; }
LOG=$OUT/sanitizer_pytests_$SAN.log
{
  echo "# pytest $ARGS with libpbx_host built $FL"
  # (leak detection off: the interpreter and torch keep memory until exit; python's own allocator is left alone)
  PBX_HOST_LIB=$TMP/libpbx_host_$SAN.so LD_PRELOAD="$PRE" \
  ASAN_OPTIONS=detect_leaks=0:abort_on_error=1:halt_on_error=1 UBSAN_OPTIONS=print_stacktrace=1 \
  TSAN_OPTIONS=halt_on_error=0:report_signal_unsafe=0:exitcode=0:log_path=$TMP/tsan_report \
    python -m pytest $ARGS -x -q -m "not gpu" -p no:cacheprovider ${KEXPR:+-k "$KEXPR"}
  echo "exit code $?"
  # reports whose stacks never enter libpbx_host are the uninstrumented interpreter / torch / OpenMP runtime
  # (their synchronisation is invisible to tsan); a report that names our library fails the run
  total=$(cat $TMP/tsan_report.* 2>/dev/null | grep -c "WARNING: ThreadSanitizer")
  ours=$(cat $TMP/tsan_report.* 2>/dev/null | grep -c "pbx_host")
  echo "thread-sanitizer reports: $total in all, $ours stack frames inside libpbx_host"
  [ "$ours" != "0" ] && cat $TMP/tsan_report.*
} > $LOG 2>&1
tail -4 $LOG
rc=1; grep -q "exit code 0" $LOG && grep -q " passed" $LOG && ! grep -q "[1-9][0-9]* stack frames inside libpbx_host" $LOG && rc=0
rm -rf $TMP
exit $rc
