#!/bin/bash
# Builds tools/bc7_host_emu/libbc7emu.so: rosy_b200/csrc/bc7_encode.cu's device functions compiled for the host
# (test tooling, see bc7_host_emu.h).  -ffp-contract=off mirrors the device build's -fmad=false.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$(dirname "$HERE")")"
CUDA_INC=${CUDA_INC:-/usr/local/cuda/include}
g++ -O2 -w -ffp-contract=off -fno-fast-math -std=c++17 -DBC7_HOST_EMU -x c++ -I"$HERE" -I"$CUDA_INC" \
    -I"$ROOT/rosy_b200/csrc" -fPIC -shared -o "$HERE/libbc7emu.so" "$ROOT/rosy_b200/csrc/bc7_encode.cu"
