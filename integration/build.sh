#!/bin/bash
# Builds integration/_build/lara_gpu: the reference's own front end (SeqAn argument parser, RNA readers, writers;
# compiled from the sources where they lie under /root/reference) with lara::solve bound to liblara_b200.so through
# the C ABI (integration/subgradient_solver_gpu.hpp).  Proof that the library is a drop-in at src/lara.cpp:57.
# Needs the reference sources, so it is built in the build container and travels to the GPU box as a binary
# (git-ignored like every other built artefact).
set -e
cd "$(dirname "$0")"
REF=${REF:-/root/reference}
[ -d "$REF/src" ] || { echo "no reference sources at $REF: keeping the prebuilt binary"; exit 0; }
mkdir -p _build
LIB=../lara_b200
if [ ! -x _build/lara_gpu ] || [ lara_gpu_main.cpp -nt _build/lara_gpu ] || [ subgradient_solver_gpu.hpp -nt _build/lara_gpu ] \
   || [ ../include/lara_b200.h -nt _build/lara_gpu ]; then
  /usr/bin/g++ -std=c++14 -O2 -DNDEBUG -DLARA_WITH_B200 -fopenmp -DWITH_OPENMP -DSEQAN_DISABLE_VERSION_CHECK -DSEQAN_HAS_ZLIB=1 \
      -isystem $REF/include/seqan/include -isystem $REF/include/lemon/include -I$REF/src -I../include -I. -w \
      lara_gpu_main.cpp -o _build/lara_gpu -L$LIB -llara_b200 -Wl,-rpath,'$ORIGIN/../../lara_b200' -lz -lpthread -lrt
fi
ls -la _build/lara_gpu
