#!/bin/bash
# Builds lara_b200/liblara_b200.so (host C++ + sm_100a CUDA, cudart linked statically so the library
# loads on a machine without a GPU) and the CLI lara_b200/lara_b200_cli.
set -e
cd "$(dirname "$0")"
OUT=${OUT:-..}      # OUT=build/dev: compile check without touching the library in the tree
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
FLAGS="-O3 -std=c++17 -lineinfo -Xcompiler -fPIC -ccbin /usr/bin/g++ --fmad=false"
B=${BUILD:-build}   # BUILD=build_x OUT=../alt EXTRA=-DLARA_BAND_DPX=1: a second build beside the one in the tree (A/B runs)
FLAGS="$FLAGS $EXTRA"
mkdir -p $B $OUT
pids=()
for f in kernels_dp kernels_band kernels_rows kernels_post kernels_setup microbench lara_gpu; do
  if [ ! -f $B/$f.o ] || [ $f.cu -nt $B/$f.o ] || [ lara_gpu.cuh -nt $B/$f.o ] || [ ../../include/lara_b200.h -nt $B/$f.o ]; then
    $NVCC $ARCH $FLAGS -I../../include -c $f.cu -o $B/$f.o &
    pids+=($!)
  fi
done
for f in lara_host lara_cli; do
  if [ ! -f $B/$f.o ] || [ host/$f.cpp -nt $B/$f.o ] || [ ../../include/lara_host.h -nt $B/$f.o ] || [ ../../include/lara_b200.h -nt $B/$f.o ]; then
    /usr/bin/g++ -O2 -std=c++17 -fPIC -I../../include -c host/$f.cpp -o $B/$f.o &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
# link into $B/ and rename: the library in the tree is always a complete file (a gpurun snapshot may be taken at any time)
$NVCC $ARCH -shared -o $B/liblara_b200.so.new $B/kernels_dp.o $B/kernels_band.o $B/kernels_rows.o $B/kernels_post.o $B/kernels_setup.o $B/microbench.o \
      $B/lara_gpu.o $B/lara_host.o $B/lara_cli.o -cudart static -ccbin /usr/bin/g++ -Xlinker --exclude-libs,ALL
mv -f $B/liblara_b200.so.new $OUT/liblara_b200.so
/usr/bin/g++ -O2 -std=c++17 -w -I../../include host/main.cpp -o $B/lara_b200_cli.new -L$OUT -llara_b200 -Wl,-rpath,'$ORIGIN' -lpthread -ldl
mv -f $B/lara_b200_cli.new $OUT/lara_b200_cli
echo built $OUT/liblara_b200.so
