#!/bin/bash
# TEST INFRASTRUCTURE: builds the UNMODIFIED reference (seqan/lara) from the sources where they lie under
# /root/reference into oracle/_ref/ (git-ignored; travels to the GPU box as a built binary).
#   lara_scalar : default CMake semantics (no -march flag -> scalar SubgradientSolver)
#   lara_simd   : -mavx2 -> SEQAN_SIMD_ENABLED -> src/subgradient_solver_simd.hpp
# The reference's own build system (cmake) is not used; this is the single g++ command of SURVEY.md 8(c).
set -e
cd "$(dirname "$0")"
REF=${REF:-/root/reference}
mkdir -p _ref
COMMON="-std=c++14 -O3 -DNDEBUG -fopenmp -DWITH_OPENMP -DLEMON_FOUND -DLEMON_ONLY_TEMPLATES -DSEQAN_DISABLE_VERSION_CHECK -DSEQAN_HAS_ZLIB=1 -isystem $REF/include/seqan/include -isystem $REF/include/lemon/include -w"
if [ ! -x _ref/lara_scalar ]; then
  /usr/bin/g++ $COMMON $REF/src/lara.cpp -o _ref/lara_scalar -lz -lpthread -lrt &
fi
if [ ! -x _ref/lara_simd ]; then
  /usr/bin/g++ $COMMON -mavx2 $REF/src/lara.cpp -o _ref/lara_simd -lz -lpthread -lrt &
fi
wait
ls -la _ref/lara_scalar _ref/lara_simd
