#!/bin/bash
# ASan + UBSan over the CPU oracle (the GPU side has compute-sanitizer in tools/sanitize.sh): builds
# oracle/librt_oracle_san.so and runs the CPU tests that drive the restatement — oracle vs the reference's host build on
# all four scenes and random scenes, golden vectors, the hand-derived known answers — with the sanitised library
# preloaded into the interpreter.  Leak checking is off (CPython itself never frees everything); any address or
# undefined-behaviour report aborts the test run.   usage: bash tools/sanitize_oracle.sh > profiles/rNN_oracle_asan.log 2>&1
set -e
cd "$(dirname "$0")/.."
make -s -C oracle librt_oracle_san.so
ASAN=$(/usr/bin/g++ -print-file-name=libasan.so); UBSAN=$(/usr/bin/g++ -print-file-name=libubsan.so)
echo "preload: $ASAN $UBSAN"
LD_PRELOAD="$ASAN $UBSAN" ASAN_OPTIONS=detect_leaks=0:abort_on_error=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1 \
  RT_ORACLE_LIB=$PWD/oracle/librt_oracle_san.so \
  python -m pytest tests/test_oracle_vs_ref.py tests/test_golden.py tests/test_zz_analytic_kat.py -q -m "not gpu" -p no:cacheprovider 2>&1 | tail -15
