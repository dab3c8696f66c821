#!/bin/bash
# Exercises the overflow-slot variant of the Zero-DCE chains (-DLLIE_OVF=1), which is not the shipped build:
# (re)builds ../libllie_ovf.so (make ovf) and runs every Zero-DCE GPU test on it through LLIE_LIB_PATH.
#   gpurun -- 'bash tools/ovf_check.sh > gpurun_out/ovf_check.txt 2>&1'
set -u
cd "$(dirname "$0")/.."
LIB=$PWD/low_light_image_enhancement_b200/libllie_ovf.so
make -C low_light_image_enhancement_b200/csrc -j8 ovf >/dev/null || exit 1   # incremental: a stale variant is rebuilt
LLIE_LIB_PATH=$LIB python -m pytest tests/test_gpu_parity.py -q -m gpu -k "dce or curve or batch or integration" "$@"
