#!/bin/bash
# Where does the colour mip kernel's time go?  Timing-only builds with the shared-memory lookups taken out
# (-DMIP_EXPERIMENT: bit 0 = no sRGB encode lookup, bit 1 = no sRGB->linear lookup; outputs are wrong on purpose).
#   tools/mip_experiment.sh build     here (no GPU): builds rosy_b200/csrc/variants/librosy_tex_cuda_mipexp{1,2,3}.so
#   tools/mip_experiment.sh           on the GPU box: tools/time_mips.py for the in-tree library and the three builds
# Result of round 2 (profiles/r2q_mip_experiment.log, 64 x 2048^2 albedo per launch): in-tree 344 us = 64 % of the HBM
# peak, without the encode table 250 us = 89 %, without the linear table (3 more ALU instructions per lookup) 303 us,
# without both 276 us.  The encode table was the bound; it was replaced by lg2 / ex2 arithmetic (272 us = 82 %).
cd "$(dirname "$0")/.."
if [ "$1" = build ]; then
    cd rosy_b200/csrc && mkdir -p variants
    NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden"
    for e in 1 2 3; do
        $NV -DMIP_EXPERIMENT=$e -DMIP_SRGB_TABLE=1 -c mip_kernels.cu -o variants/mip_$e.o
        $NV -shared -o variants/librosy_tex_cuda_mipexp$e.so bc7_encode.o bc45_encode.o variants/mip_$e.o rosy_tex_cuda.o image_io.o jpeg_decode.o -lz -cudart static
        rm -f variants/mip_$e.o
    done
    exit 0
fi
for v in main mipexp1 mipexp2 mipexp3; do
  if [ "$v" = main ]; then unset ROSY_TEX_LIB; else export ROSY_TEX_LIB=$PWD/rosy_b200/csrc/variants/librosy_tex_cuda_$v.so; fi
  echo "== $v"; python tools/time_mips.py 2>&1 | grep -E '"albedo"' | grep -E '"textures": 64|"size": 2048, "textures": 1,'
done
