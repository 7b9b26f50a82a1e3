#!/bin/bash
# tools/build_variant.sh <tag> <nvcc -D flags...>: an A/B build of libsparkle_cuda next to the default one
# (sparkle_b200/libsparkle_cuda_<tag>.so; selected with SPK_CUDA_LIB=libsparkle_cuda_<tag>.so). Kernel experiments only.
set -e
cd "$(dirname "$0")/../sparkle_b200"
tag=$1; shift
mkdir -p build_$tag
NVFLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v"
objs=""
for f in spk_api spk_trace spk_render spk_wavefront spk_denoise spk_group spk_build; do
  case $f in
    spk_trace|spk_wavefront|spk_render) /usr/local/cuda/bin/nvcc $NVFLAGS "$@" -c csrc/$f.cu -o build_$tag/$f.o 2> build_$tag/$f.ptxas.log || { cat build_$tag/$f.ptxas.log; exit 1; }
       objs="$objs build_$tag/$f.o";;
    *) objs="$objs build/$f.o";;
  esac
done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o libsparkle_cuda_$tag.so $objs -cudart static -Xlinker --exclude-libs,ALL
grep -A2 "wf_extend\|trace_persistent" build_$tag/spk_wavefront.ptxas.log build_$tag/spk_trace.ptxas.log | grep "Used\|spill" | sort | uniq -c | sort -rn | head -8
