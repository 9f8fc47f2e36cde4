#!/bin/bash
# A/B builds: tools/build_variant.sh <name> <file.cu> "<extra nvcc flags>"  ->  variants/libvfi_<name>.so
# (every other object is reused from the normal build; select at run time with VFI_B200_LIB=<path>)
set -eu
NAME=$1; SRC=$2; EXTRA=${3:-}
python -m videoframeinterpolation_b200.build > /dev/null
mkdir -p variants
OBJ=videoframeinterpolation_b200/build
STEM=$(basename $SRC .cu)
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr $EXTRA -c $SRC -o variants/${STEM}_${NAME}.o
OBJS=$(ls $OBJ/*.o | grep -v "/${STEM}.o")
nvcc -shared -o variants/libvfi_${NAME}.so $OBJS variants/${STEM}_${NAME}.o -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -lcudart_static -lcuda
echo variants/libvfi_${NAME}.so
