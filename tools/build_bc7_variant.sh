#!/bin/bash
# Build a variant of the library with extra -D flags for the BC7 kernel (A/B timing aid):
#   tools/build_bc7_variant.sh sync512 "-DBC7_NT=512 -DBC7_JOB_SYNC=1"
# -> rosy_b200/csrc/variants/librosy_tex_cuda_<name>.so ; run with ROSY_TEX_LIB=<that path>
set -e
NAME=$1; FLAGS=$2
cd "$(dirname "$0")/../rosy_b200/csrc"
mkdir -p variants
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden"
$NV -fmad=false -prec-div=true -prec-sqrt=true $FLAGS -c bc7_encode.cu -o variants/bc7_$NAME.o
$NV -shared -o variants/librosy_tex_cuda_$NAME.so variants/bc7_$NAME.o bc45_encode.o mip_kernels.o rosy_tex_cuda.o image_io.o jpeg_decode.o -lz -cudart static
rm -f variants/bc7_$NAME.o
echo built variants/librosy_tex_cuda_$NAME.so
