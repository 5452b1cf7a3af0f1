#!/bin/bash
# Build an A/B variant of the library: scripts/build_variant.sh NAME -DFOO=1 ...  → build/libminipath_b200_NAME.so (load with MINIPATH_B200_LIB)
set -e
name=$1; shift
mkdir -p build
CS=minipath_b200/csrc
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -Xcompiler -fPIC,-ffp-contract=off -Xptxas -v "$@" -c $CS/context.cu -o build/context_$name.o 2> build/ptxas_$name.log
/usr/local/cuda/bin/nvcc -shared -o build/libminipath_b200_$name.so $CS/host_bvh.o $CS/host_camera.o $CS/host_render.o $CS/host_image.o $CS/host_group.o $CS/build_device.o build/context_$name.o -lpthread -lrt -ldl -cudart static 2>/dev/null
echo built build/libminipath_b200_$name.so
