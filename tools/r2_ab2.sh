#!/bin/bash
# Developer tool (GPU box): fill on a high-priority stream / smaller fill blocks / workers, at 1/8 share and on all of C3
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "== parity FILLPRIO=1 FILLWARPS=1 WORKERS=3"
LARA_B200_FILLPRIO=1 LARA_B200_FILLWARPS=1 python -m pytest tests/test_gpu_solve.py tests/test_gpu_band.py -m gpu -x -q 2>&1 | tail -n 2
echo "== sweep 1/8 share"
NSEQ=500 STEPS=2 EXTRA="--emulate-world 8" tools/sweep_env.sh "" "LARA_B200_FILLWARPS=1" "LARA_B200_FILLPRIO=1" "LARA_B200_FILLPRIO=1 LARA_B200_FILLWARPS=1" "LARA_B200_FILLPRIO=1 LARA_B200_FILLWARPS=1 LARA_B200_WORKERS=3" "LARA_B200_FILLPRIO=1 LARA_B200_FILLWARPS=1 LARA_B200_WORKERS=4" "LARA_B200_FILLPRIO=1 LARA_B200_FILLWARPS=2 LARA_B200_WORKERS=3" "LARA_B200_FILLWARPS=1 LARA_B200_WORKERS=3"
echo "== sweep N=1"
NSEQ=500 STEPS=2 tools/sweep_env.sh "LARA_B200_FILLPRIO=1 LARA_B200_FILLWARPS=1" "LARA_B200_FILLWARPS=1" "LARA_B200_FILLPRIO=1 LARA_B200_FILLWARPS=1 LARA_B200_WORKERS=3"
} > gpurun_out/r2_ab2.log 2>&1
tail -n 30 gpurun_out/r2_ab2.log
