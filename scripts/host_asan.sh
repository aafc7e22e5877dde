#!/bin/bash
# -fsanitize=address,undefined build of the host front end (loader, writers, C surface) under the CPU
# tests that exercise it.  The sanitised librtm_host.so replaces the normal one for the run and is removed afterwards.
set -e
cd "$(dirname "$0")/.."
H=raytracingmin_b200/host
CXX=${CXX:-g++}
cp $H/librtm_host.so /tmp/librtm_host.keep.so
$CXX -std=c++17 -O1 -g -fno-omit-frame-pointer -fsanitize=address,undefined -fno-sanitize-recover=undefined -fPIC -ffp-contract=off \
    -shared -o $H/librtm_host.so $H/host_capi.cpp $H/SettingData.cpp
ASAN=$($CXX -print-file-name=libasan.so)
set +e
LD_PRELOAD=$ASAN ASAN_OPTIONS=detect_leaks=0:abort_on_error=1 UBSAN_OPTIONS=print_stacktrace=1 \
    python -m pytest tests/test_host_frontend.py tests/test_planes_host.py -q -m "not gpu" 2>&1 | tail -15
rc=${PIPESTATUS[0]}
cp /tmp/librtm_host.keep.so $H/librtm_host.so
echo "host asan/ubsan rc=$rc"
exit $rc
