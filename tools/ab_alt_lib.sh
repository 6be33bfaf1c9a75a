#!/bin/bash
# Developer tool (GPU box): A/B of a second build (lara_b200/alt/liblara_b200.so, made with
#   BUILD=build_x OUT=../alt EXTRA=-D... bash lara_b200/csrc/build.sh) against the library in the tree: parity, 1/8 share, all of C3
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
ALT=$PWD/lara_b200/alt/liblara_b200.so
{
echo "== parity alt"
LARA_B200_LIB=$ALT python -m pytest tests/test_gpu_band.py tests/test_gpu_solve.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -n 2
echo "== 1/8 share"
NSEQ=500 STEPS=3 EXTRA="--emulate-world 8" tools/sweep_env.sh "" "LARA_B200_LIB=$ALT" "" "LARA_B200_LIB=$ALT"
echo "== all of C3"
NSEQ=500 STEPS=2 tools/sweep_env.sh "" "LARA_B200_LIB=$ALT"
} > gpurun_out/r2_ab4.log 2>&1
tail -n 14 gpurun_out/r2_ab4.log
