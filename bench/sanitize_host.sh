#!/bin/bash
# ASan + UBSan and TSan builds of the C++ host runtime, run under the host-native test files
# (the reference has no sanitizer settings; this is the B200-side analogue for the
# scheduler / channel / record code).  Usage: bench/sanitize_host.sh [asan|tsan]
set -e
cd "$(dirname "$0")/.."
mode=${1:-asan}
out=/tmp/dryad_san_$mode
mkdir -p $out
if [ "$mode" = tsan ]; then flags="-fsanitize=thread"; rt=$(g++ -print-file-name=libtsan.so);
else flags="-fsanitize=address,undefined -fno-sanitize-recover=undefined"; rt=$(g++ -print-file-name=libasan.so); fi
g++ -std=c++17 -O1 -g $flags -fno-omit-frame-pointer -fPIC -shared -Iinclude csrc/host/*.cpp -o $out/libdryad_host.so -lz -lpthread
ASAN_OPTIONS=detect_leaks=0:halt_on_error=1 TSAN_OPTIONS=halt_on_error=1:report_signal_unsafe=0 LD_PRELOAD=$rt DRYAD_HOST_LIB=$out/libdryad_host.so \
  python -m pytest tests/test_host_native.py tests/test_fault_tolerance.py tests/test_services.py -x -q -p no:cacheprovider 2>&1 | tail -15
