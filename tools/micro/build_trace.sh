#!/bin/bash
# Builds build_micro/ring_trace: the kernel library compiled with -DLLIE_RING_TRACE plus the timeline harness.
set -e
cd "$(dirname "$0")/../.."
mkdir -p build_micro
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -prec-div=true -prec-sqrt=true -DLLIE_RING_TRACE -Iinclude"
nvcc $FLAGS $EXTRA -o build_micro/${OUT:-ring_trace} tools/micro/ring_trace.cu lowlight_image_enhancement_b200/csrc/*.cu
