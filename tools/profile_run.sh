#!/bin/bash
# Developer tool (GPU box): one profiling pass over a one-batch workload (first NSEQ sequences of C3, one worker):
#   1. plain bench run (must exit 0 before anything runs under ncu)
#   2. ncu launch list (gpu__time_duration.sum) -> gpurun_out/$TAG_launches.csv
#   3. ncu --set full of a few launches of the DP and post kernels -> gpurun_out/$TAG_full.ncu-rep
TAG=${TAG:-prof}
NSEQ=${NSEQ:-100}
export LARA_B200_WORKERS=1
CMD="python bench.py --nseq $NSEQ --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || { echo "bench failed"; tail -5 gpurun_out/${TAG}_bench.err; exit 1; }
tail -2 gpurun_out/${TAG}_bench.err; cat gpurun_out/${TAG}_bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -s ${SKIP:-600} -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-gotoh_|iterate_post}" -s ${SKIPFULL:-300} -c ${NFULL:-6} -o gpurun_out/${TAG}_full -f $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ls -la gpurun_out/ | grep ${TAG}
