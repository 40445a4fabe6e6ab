#!/bin/bash
# scripts/build_variant.sh NAME "-DFLAG=..."  -> variants/libcanvas_b200.NAME.so (tuning experiments; not shipped)
set -e
name=$1; flags=$2
cd "$(dirname "$0")/../canvas_b200/csrc"
mkdir -p ../../variants build
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -Xcompiler -fPIC,-fvisibility=default \
    -I../../include -I/usr/local/cuda/include $flags -Xptxas -v -c cv_kernels.cu -o build/cv_kernels.$name.o 2> build/ptxas.$name.log
objs=$(ls build/*.o | grep -v "cv_kernels")
/usr/local/cuda/bin/nvcc -shared -o ../../variants/libcanvas_b200.$name.so $objs build/cv_kernels.$name.o -lpthread -lm
rm -f build/cv_kernels.$name.o
echo built variants/libcanvas_b200.$name.so
