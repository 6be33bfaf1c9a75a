#!/bin/bash
# Developer tool (GPU box): one profiling pass over the bench workload (C3, or its first NSEQ sequences):
#   1. plain bench run (must exit 0 before anything runs under ncu)
#   2. ncu launch list (gpu__time_duration.sum) of the same command -> gpurun_out/${TAG}_launches.csv
#   3. ncu --set full of a few launches of the DP and post kernels (single worker: launches back to back)
#      -> gpurun_out/${TAG}_full.ncu-rep
TAG=${TAG:-prof}
ARGS="--steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-isolated --no-traffic"
[ -n "$NSEQ" ] && ARGS="--nseq $NSEQ $ARGS"
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || { echo "bench failed"; tail -5 gpurun_out/${TAG}_bench.err; exit 1; }
tail -1 gpurun_out/${TAG}_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -s ${SKIP:-600} -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py $ARGS > gpurun_out/${TAG}_ncu1.log 2>&1
LARA_B200_WORKERS=1 ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-gotoh_|iterate_post}" -s ${SKIPFULL:-60} -c ${NFULL:-4} -o gpurun_out/${TAG}_full -f python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-isolated --no-traffic ${NSEQ:+--nseq $NSEQ} > gpurun_out/${TAG}_ncu2.log 2>&1
ls -la gpurun_out/ | grep ${TAG}
