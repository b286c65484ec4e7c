#!/bin/bash
# Build an experimental copy of the library with extra nvcc flags, for A/B timing on the GPU box:
#   bash tools/build_exp.sh NAME -DRT_SOMETHING ...   ->  tools/_build/librtcuda_NAME.so
#   RTCUDA_LIB=tools/_build/librtcuda_NAME.so python tools/time_variants.py 1920 1080 64 3
set -e
NAME=$1; shift
cd "$(dirname "$0")/../gpuraytracer_b200/csrc"
B=../../tools/_build; mkdir -p $B
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo --fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
    "$@" -cubin -o $B/pt_$NAME.cubin pathtrace.cu
/usr/local/cuda/bin/bin2c --const --name rt_cubin_image $B/pt_$NAME.cubin > $B/pt_${NAME}_cubin.c
echo 'const unsigned long long rt_cubin_image_size = sizeof(rt_cubin_image);' >> $B/pt_${NAME}_cubin.c
# (build/noalloc.o: the runtime-API translation unit of the main build, make -C gpuraytracer_b200/csrc first)
g++ -O2 -std=c++17 -fPIC -ffp-contract=off -fno-fast-math -I/usr/local/cuda/include -pthread -shared -o $B/librtcuda_$NAME.so rtcuda.cpp rt_plan.cpp host_scene.cpp host_bvh.cpp host_renderer.cpp build/noalloc.o -x c $B/pt_${NAME}_cubin.c \
    -L/usr/local/cuda/lib64 -lcudart_static -Wl,--exclude-libs,libcudart_static.a -ldl -lrt -lpthread
echo $B/librtcuda_$NAME.so
