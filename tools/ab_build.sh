#!/bin/bash
# Build a variant of the library for A/B timing: tools/ab_build.sh NAME "-DFOO=1 ..."  -> ab/libwebpb200_NAME.so
# (run with WEBPB200_SO=ab/libwebpb200_NAME.so python bench.py ...)
set -e
cd "$(dirname "$0")/.."
mkdir -p ab
make -s -C image-webp_b200/csrc BUILD=build_$1 OUT=../../ab/libwebpb200_$1.so EXTRA="$2"
grep -E "Used" image-webp_b200/csrc/build_$1/kernels.ptxas.log | tail -2
