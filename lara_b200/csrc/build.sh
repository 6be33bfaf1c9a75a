#!/bin/bash
# Builds lara_b200/liblara_b200.so (host C++ + sm_100a CUDA, cudart linked statically so the library
# loads on a machine without a GPU) and the CLI lara_b200/lara_b200_cli.
set -e
cd "$(dirname "$0")"
OUT=${OUT:-..}      # OUT=build/dev: compile check without touching the library in the tree
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
FLAGS="-O3 -std=c++17 -lineinfo -Xcompiler -fPIC -ccbin /usr/bin/g++ --fmad=false"
mkdir -p build $OUT
pids=()
for f in kernels_dp kernels_band kernels_rows kernels_post kernels_setup microbench lara_gpu; do
  if [ ! -f build/$f.o ] || [ $f.cu -nt build/$f.o ] || [ lara_gpu.cuh -nt build/$f.o ] || [ ../../include/lara_b200.h -nt build/$f.o ]; then
    $NVCC $ARCH $FLAGS -I../../include -c $f.cu -o build/$f.o &
    pids+=($!)
  fi
done
for f in lara_host lara_cli; do
  if [ ! -f build/$f.o ] || [ host/$f.cpp -nt build/$f.o ] || [ ../../include/lara_host.h -nt build/$f.o ] || [ ../../include/lara_b200.h -nt build/$f.o ]; then
    /usr/bin/g++ -O2 -std=c++17 -fPIC -I../../include -c host/$f.cpp -o build/$f.o &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
# link into build/ and rename: the library in the tree is always a complete file (a gpurun snapshot may be taken at any time)
$NVCC $ARCH -shared -o build/liblara_b200.so.new build/kernels_dp.o build/kernels_band.o build/kernels_rows.o build/kernels_post.o build/kernels_setup.o build/microbench.o \
      build/lara_gpu.o build/lara_host.o build/lara_cli.o -cudart static -ccbin /usr/bin/g++ -Xlinker --exclude-libs,ALL
mv -f build/liblara_b200.so.new $OUT/liblara_b200.so
/usr/bin/g++ -O2 -std=c++17 -w -I../../include host/main.cpp -o build/lara_b200_cli.new -L$OUT -llara_b200 -Wl,-rpath,'$ORIGIN' -lpthread -ldl
mv -f build/lara_b200_cli.new $OUT/lara_b200_cli
echo built $OUT/liblara_b200.so
