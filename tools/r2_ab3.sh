#!/bin/bash
# Developer tool (GPU box): parity + speed of the current library on C3 (whole, 1/8 share), C5-400, C4
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "== parity"
python -m pytest tests/test_gpu_band.py tests/test_gpu_solve.py tests/test_gpu_golden.py tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -n 3
echo "== speed"
NSEQ=500 STEPS=2 EXTRA="--emulate-world 8" tools/sweep_env.sh "" $SWEEP8
NSEQ=500 STEPS=2 tools/sweep_env.sh "" $SWEEP1
NSEQ=400 STEPS=1 EXTRA="--workload c5" tools/sweep_env.sh ""
NSEQ=50 STEPS=2 EXTRA="--workload c4" tools/sweep_env.sh ""
} > gpurun_out/r2_ab3.log 2>&1
tail -n 30 gpurun_out/r2_ab3.log
