#!/bin/bash
# Development aid: rebuild only the tendency translation unit(s) with extra -D switches and relink, e.g.
#   tools/variant.sh "-DTF_MINBLOCKS=3" && python tools/time_phases.py 256 5 fma
#   tools/variant.sh "-DTF_ROTATE_EXACT=1" exact && python tools/time_phases.py 256 5
# (run on the GPU box inside one gpurun call to compare kernel variants without a full rebuild)
set -e
cd "$(dirname "$0")/../oceananigans.jl_b200/csrc"
ARCH="-gencode arch=compute_100a,code=sm_100a"
FLAGS="-O3 -std=c++17 $ARCH -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v"
if [ "$2" = "exact" ]; then
  nvcc $FLAGS -fmad=false $1 -c model.cu -o build/model.o 2> build/model.ptxas.log
  grep -A2 "tendency_fast_kernelIdLi0\|tendency_tiled_kernelIdLi0" build/model.ptxas.log | grep -E "registers|spill" | head -3
else
  nvcc $FLAGS -fmad=true $1 -c tendency_fma.cu -o build/tendency_fma.o 2> build/tendency_fma.ptxas.log
  grep -A2 "tendency_plane_kernelIdLi1\|tendency_fast_kernelIdLi1" build/tendency_fma.ptxas.log | grep -E "registers|spill" | head -3
fi
nvcc $ARCH -shared -o ../libnhstep.so build/api.o build/model.o build/dist.o build/tendency_fma.o build/solver_generic.o \
     build/no_alloc.o build/grid_host.o -cudart static
