#!/bin/bash
# The whole `-m gpu` suite against the checks build (libodk_checks.so: device-side bounds assertions, ODK_ASSERT in
# csrc/odk_common.cuh) — the stand-in for compute-sanitizer, which is closed on this GPU pool.  A failed assertion traps
# the kernel and fails the test with a CUDA error.   Usage: gpurun -- 'bash scripts/gpu_checks.sh'
mkdir -p gpurun_out
LOG=gpurun_out/checks_build.log
echo "# ODK_LIB_VARIANT=checks python -m pytest tests -m gpu -q   (libodk_checks.so = same sources, -DODK_CHECKS=1)" > $LOG
ODK_LIB_VARIANT=checks timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider >> $LOG 2>&1
echo "exit $?" >> $LOG
echo "ODK_ASSERT sites in the sources: $(cat object_detector_b200/csrc/*.cu | grep -c ODK_ASSERT)" >> $LOG
tail -6 $LOG
