#!/bin/bash
# Experimental build of libpltb200 with extra nvcc flags:  tools/build_variant.sh <suffix> <flags...>  ->  mitsuba3_plt_b200/libpltb200_<suffix>.so
# (selected at run time with PLTB200_LIB, see tools/gpu.sh ab:<suffix>; variants are not tracked by git)
set -e
sfx=$1; shift
root=$(cd $(dirname $0)/.. && pwd)
src=$root/mitsuba3_plt_b200/csrc
tmp=$(mktemp -d)
# EXACT_RSQRT=1 drops -DPLTB_FAST_RSQRT (nvcc applies -U before -D, so it cannot be undone with a flag)
base="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC"
[ -z "$EXACT_RSQRT" ] && base="$base -DPLTB_FAST_RSQRT"
nvcc $base "$@" -Xptxas -v -c -o $tmp/k.o $src/pltb_kernels.cu 2> $root/gpurun_out/ptxas_$sfx.log
nvcc $base "$@" -c -o $tmp/c.o $src/pltb_capi.cu
nvcc $base "$@" -c -o $tmp/n.o $src/pltb_noalloc.cu
nvcc -shared -o $root/mitsuba3_plt_b200/libpltb200_$sfx.so $tmp/k.o $tmp/c.o $tmp/n.o -ldl 2>/dev/null
rm -rf $tmp
echo built libpltb200_$sfx.so
