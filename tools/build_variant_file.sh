#!/bin/bash
# tools/build_variant_file.sh NAME FILE.cu "-DMACRO=..."  ->  randblas_b200/variants/librbx_NAME.so (only FILE.cu is recompiled with the macros)
set -e
HERE=$(cd "$(dirname "$0")/.." && pwd)
C=$HERE/randblas_b200/csrc
mkdir -p $HERE/randblas_b200/variants $C/_obj
make -C $C -j8 >/dev/null
B=$(basename $2 .cu)
F="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-O2 --expt-relaxed-constexpr -ccbin /usr/bin/g++"
/usr/local/cuda/bin/nvcc $F $3 -c $C/$B.cu -o $C/_obj/${B}_$1_var.o
OBJS=$(ls $C/_obj/*.o | grep -v "/$B.o" | grep -v "_var.o")
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $HERE/randblas_b200/variants/librbx_$1.so $OBJS $C/_obj/${B}_$1_var.o -ldl
echo built $1
