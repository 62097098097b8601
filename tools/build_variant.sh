#!/bin/bash
# tools/build_variant.sh NAME "-DMACRO=..."  ->  randblas_b200/variants/librbx_NAME.so (only rbx_saso.cu is recompiled with the macros)
set -e
HERE=$(cd "$(dirname "$0")/.." && pwd)
C=$HERE/randblas_b200/csrc
mkdir -p $HERE/randblas_b200/variants $C/_obj
make -C $C -j8 >/dev/null
F="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-O2 --expt-relaxed-constexpr -ccbin /usr/bin/g++"
/usr/local/cuda/bin/nvcc $F $2 -c $C/rbx_saso.cu -o $C/_obj/rbx_saso_$1.o
OBJS=$(ls $C/_obj/*.o | grep -v "rbx_saso")
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $HERE/randblas_b200/variants/librbx_$1.so $OBJS $C/_obj/rbx_saso_$1.o -ldl
echo built $1
