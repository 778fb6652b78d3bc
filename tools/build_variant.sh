#!/bin/bash
# developer helper: build a probe variant of the library next to the shipped one.  tools/build_variant.sh <tag> <extra nvcc flags...>
tag=$1; shift
cd "$(dirname "$0")/.."
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false --extended-lambda -shared -Xcompiler -fPIC "$@" -o mazetd_b200/variants/libmazetd_b200_$tag.so mazetd_b200/csrc/capi.cu && echo "built $tag"
