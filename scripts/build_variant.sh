#!/bin/bash
# scripts/build_variant.sh NAME "-DFLAG=..." [FILE.cu]  -> variants/libcanvas_b200.NAME.so (tuning experiments; not shipped)
# FILE.cu is the kernel translation unit the flags apply to (default cv_kernels.cu; cv_rect.cu compiles in 15 s).
set -e
name=$1; flags=$2; unit=${3:-cv_kernels.cu}
base=${unit%.cu}
cd "$(dirname "$0")/../canvas_b200/csrc"
mkdir -p ../../variants build
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -Xcompiler -fPIC \
    -I../../include -I/usr/local/cuda/include $flags -Xptxas -v -c $unit -o build/variant.$name.o 2> build/ptxas.$name.log
objs=$(ls build/*.o | grep -v "build/$base.o" | grep -v "build/variant\.")
/usr/local/cuda/bin/nvcc -shared -o ../../variants/libcanvas_b200.$name.so $objs build/variant.$name.o -Xlinker --version-script=exports.map -lpthread -lm
rm -f build/variant.$name.o
echo built variants/libcanvas_b200.$name.so
