for v in main mipexp1 mipexp2 mipexp3; do
  if [ "$v" = main ]; then unset ROSY_TEX_LIB; else export ROSY_TEX_LIB=$PWD/rosy_b200/csrc/variants/librosy_tex_cuda_$v.so; fi
  echo "== $v"; python tools/time_mips.py 2>&1 | grep -E '"albedo"' | grep -E '"textures": 64|"size": 2048, "textures": 1,'
done
