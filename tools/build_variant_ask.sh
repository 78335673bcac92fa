#!/bin/bash
# usage: tools/build_variant_ask.sh <suffix> <extra nvcc flags...>  -> trackmaker-rs_b200/libtmkphy_<suffix>.so (ask.cu rebuilt with the flags)
suf=$1; shift
cd "$(dirname "$0")/.."
C=trackmaker-rs_b200/csrc
mkdir -p $C/build/var_$suf
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -O2 "$@" -c $C/ask.cu -o $C/build/var_$suf/ask.o || exit 1
objs=$(ls $C/build/*.o | grep -v "/ask.o")
/usr/local/cuda/bin/nvcc -shared -o trackmaker-rs_b200/libtmkphy_$suf.so $objs $C/build/var_$suf/ask.o -gencode arch=compute_100a,code=sm_100a -lcudart
