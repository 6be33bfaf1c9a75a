#!/bin/bash
# Developer tool (GPU box): parity of the alternative builds / switches, then an A/B sweep on C3 (whole and 1/8 share).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
ALT=$PWD/lara_b200/alt/liblara_b200.so
{
echo "== parity: alt library (LARA_BAND_DPX=1)"
LARA_B200_LIB=$ALT python -m pytest tests/test_gpu_band.py tests/test_gpu_solve.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -n 3
echo "== parity: POSTREGS=64 SMEMCAND=128"
LARA_B200_POSTREGS=64 LARA_B200_SMEMCAND=128 python -m pytest tests/test_gpu_solve.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -n 3
echo "== sweep N=1"
NSEQ=500 STEPS=2 tools/sweep_env.sh "" "LARA_B200_LIB=$ALT" "LARA_B200_POSTREGS=64 LARA_B200_SMEMCAND=128" "LARA_B200_SMEMCAND=128" "LARA_B200_LIB=$ALT LARA_B200_POSTREGS=64 LARA_B200_SMEMCAND=128"
echo "== sweep 1/8 share"
NSEQ=500 STEPS=3 EXTRA="--emulate-world 8" tools/sweep_env.sh "" "LARA_B200_LIB=$ALT" "LARA_B200_POSTREGS=64 LARA_B200_SMEMCAND=128" "LARA_B200_LIB=$ALT LARA_B200_POSTREGS=64 LARA_B200_SMEMCAND=128"
} > gpurun_out/r2_ab.log 2>&1
tail -n 30 gpurun_out/r2_ab.log
