#!/bin/bash
# AddressSanitizer + UBSan over the host side of libllie (everything in the .cu files that is not device code).
# Builds ../libllie_asan.so (make asan) and runs pytest on it through LLIE_LIB_PATH with the sanitizer runtimes
# preloaded into python.  Without a GPU (this container):   bash tools/host_sanitize.sh -m "not gpu"
# On a GPU box (kernels run uninstrumented, the host code around them is checked):
#   gpurun -- 'bash tools/host_sanitize.sh -m gpu -k "sharded or tiled or stream or no_alloc or errors" > gpurun_out/asan_gpu.txt 2>&1'
set -u
cd "$(dirname "$0")/.."
make -C low_light_image_enhancement_b200/csrc -j8 asan >/dev/null || exit 1
export LLIE_LIB_PATH=$PWD/low_light_image_enhancement_b200/libllie_asan.so
export LD_PRELOAD="$(gcc -print-file-name=libasan.so):$(gcc -print-file-name=libubsan.so)"
# python itself leaks by design; CUDA maps memory inside ASan's shadow gap
export ASAN_OPTIONS=detect_leaks=0:protect_shadow_gap=0:halt_on_error=1:abort_on_error=0
export UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1
python -m pytest tests/test_abi.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py tests/test_sharding_gloo.py -q -x "$@"
