#!/bin/bash
# Host runtime under AddressSanitizer + UBSan and under ThreadSanitizer (the GPU pool refuses
# compute-sanitizer, the host side has no such limit).  Links csrc/host/*.cc straight into the driver.
#   benchmarks/sanitize_host.sh [outdir]        -> <outdir>/sanitizer_host_{asan,tsan}.log
set -u
cd "$(dirname "$0")/.."
OUT=${1:-profiles}
TMP=$(mktemp -d)
SRC="csrc/host/*.cc csrc/host/tests/sanitize_driver.cc"
rc=0
for mode in asan tsan; do
  if [ $mode = asan ]; then FL="-fsanitize=address,undefined -fno-sanitize-recover=undefined"; else FL="-fsanitize=thread"; fi
  g++ -std=c++17 -O1 -g -fno-omit-frame-pointer $FL -pthread -o $TMP/drv_$mode $SRC -ldl 2> $TMP/build_$mode.log || { echo "build failed ($mode)"; cat $TMP/build_$mode.log; rc=1; continue; }
  mkdir -p $TMP/run_$mode
  { echo "# $mode: g++ $FL"; ASAN_OPTIONS=detect_leaks=1 TSAN_OPTIONS=halt_on_error=0 $TMP/drv_$mode $TMP/run_$mode; echo "exit code $?"; } > $OUT/sanitizer_host_$mode.log 2>&1
  tail -2 $OUT/sanitizer_host_$mode.log
  grep -q "exit code 0" $OUT/sanitizer_host_$mode.log || rc=1
done
rm -rf $TMP
exit $rc
