#!/bin/bash
# Builds blitzdg_b200/csrc/libblitzdg_b200_<name>.so from the product sources with extra -D flags
# (kernel experiments, -DBDG_PHASEPROF).  Select it at run time with BDG_B200_LIB=<path>.
#   tools/build_variant.sh prof -DBDG_PHASEPROF
set -e
name=$1; shift
here=$(cd "$(dirname "$0")/.." && pwd)
src=$here/blitzdg_b200/csrc
obj=$here/build/variant_$name
mkdir -p $obj
nccl=$(python -c "import nvidia.nccl; print(list(nvidia.nccl.__path__)[0])")
for f in capi sw2d_generic sw2d_fused sw2d_halo sw2d_order sw2d_volume_dmma sw2d_output sw2d_tracer sw2d_noalloc sw2d_curved; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -I$nccl/include \
    -Xptxas -v "$@" -c $src/$f.cu -o $obj/$f.o 2> $obj/$f.ptxas.log &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $src/libblitzdg_b200_$name.so $obj/*.o \
  -L$nccl/lib -l:libnccl.so.2 -Xlinker -rpath -Xlinker $nccl/lib
grep -A2 "k_fused_persistentILi15" $obj/sw2d_fused.ptxas.log | grep -E "registers|spill" || true
echo built $src/libblitzdg_b200_$name.so
