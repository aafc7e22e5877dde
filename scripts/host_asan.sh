#!/bin/bash
# -fsanitize=address,undefined build of the host front end (loader, JSON reader/writer, BMP/JPEG
# writers, C surface) driven by tests/native/host_sanitize_driver.cpp.  Log: profiles/r02/host_asan_ubsan.log
cd "$(dirname "$0")/.."
H=raytracingmin_b200/host
CXX=${CXX:-/usr/bin/g++}
T=$(mktemp -d)
$CXX -std=c++17 -O1 -g -fno-omit-frame-pointer -fsanitize=address,undefined -fno-sanitize-recover=undefined -ffp-contract=off \
    -o $T/host_san tests/native/host_sanitize_driver.cpp $H/host_capi.cpp $H/SettingData.cpp -B/usr/lib/gcc/x86_64-linux-gnu/13 -L/usr/lib/gcc/x86_64-linux-gnu/13 || exit 2
ASAN_OPTIONS=detect_leaks=1 UBSAN_OPTIONS=print_stacktrace=1 $T/host_san scenes $T
rc=$?
echo "host asan/ubsan rc=$rc"
rm -rf $T
exit $rc
