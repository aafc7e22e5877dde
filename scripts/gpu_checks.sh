#!/bin/bash
# Memory-safety evidence without compute-sanitizer (closed on this pool): the GPU test-suite against the library built with
# the device-side invariants compiled in (make -C raytracingmin_b200/csrc checks).  A violated invariant = cudaErrorAssert.
mkdir -p gpurun_out
[ -f raytracingmin_b200/librtm_b200_checks.so ] || { echo "build it first: make -C raytracingmin_b200/csrc checks"; exit 2; }
cp raytracingmin_b200/librtm_b200.so /tmp/librtm_keep.so
cp raytracingmin_b200/librtm_b200_checks.so raytracingmin_b200/librtm_b200.so
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/device_checks_pytest.log 2>&1; rc=$?
cp /tmp/librtm_keep.so raytracingmin_b200/librtm_b200.so
echo "device-checks build under the GPU tests: rc=$rc"; tail -4 gpurun_out/device_checks_pytest.log
