#!/bin/bash
# A/B of BC7 kernel variants built by tools/build_bc7_variant.sh: per variant the time of one 2048^2 chain's block
# launch (CUDA events, tools/ab_bc7.py) and its warp-instruction count (ncu, one launch).  Run on the GPU box:
#   tools/ab_bc7_variants.sh none keep konst ...   ("main" = the in-tree library)
OUT=gpurun_out/ab_variants.log
: > $OUT
for v in "$@"; do
    if [ "$v" = main ]; then unset ROSY_TEX_LIB; else export ROSY_TEX_LIB=$PWD/rosy_b200/csrc/variants/librosy_tex_cuda_$v.so; fi
    t=$(python tools/ab_bc7.py one 2>&1 | tail -1)
    ncu --metrics smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio -k regex:bc7_encode -s 3 -c 1 --csv python tools/ab_bc7.py one > gpurun_out/ab_ncu_$v.csv 2>&1
    i=$(grep "smsp__inst_executed.sum" gpurun_out/ab_ncu_$v.csv | tail -1 | awk -F'","' '{print $NF}' | tr -d '",')
    l=$(grep "thread_inst_executed_per_inst" gpurun_out/ab_ncu_$v.csv | tail -1 | awk -F'","' '{print $NF}' | tr -d '",')
    echo "$v | $t | warp-inst $i | lanes $l" | tee -a $OUT
done
