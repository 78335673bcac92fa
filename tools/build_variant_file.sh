#!/bin/bash
# usage: tools/build_variant_file.sh <file stem> <suffix> <extra nvcc flags...>  -> trackmaker-rs_b200/libtmkphy_<suffix>.so (csrc/<stem>.cu rebuilt with the flags)
stem=$1; suf=$2; shift 2
cd "$(dirname "$0")/.."
C=trackmaker-rs_b200/csrc
mkdir -p $C/build/var_$suf
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -O2 "$@" -c $C/$stem.cu -o $C/build/var_$suf/$stem.o || exit 1
objs=$(ls $C/build/*.o | grep -v "/$stem.o")
/usr/local/cuda/bin/nvcc -shared -o trackmaker-rs_b200/libtmkphy_$suf.so $objs $C/build/var_$suf/$stem.o -gencode arch=compute_100a,code=sm_100a -lcudart
