#!/bin/bash
# Developer tool (GPU box): bench.py over LARA_B200_WORKERS : LARA_B200_BATCH : LARA_B200_BANDQ (0 = automatic);
# prints one short line per run.   usage: NSEQ=500 tools/sweep.sh 3:8192:0 1:65536:16 ...
NSEQ=${NSEQ:-200}
for cfg in "$@"; do
  IFS=: read W B Q <<< "$cfg"
  LARA_B200_WORKERS=$W LARA_B200_BATCH=$B LARA_B200_BANDQ=${Q:-0} timeout 900 python bench.py --nseq $NSEQ --steps ${STEPS:-1} --warmup 1 --no-cpu-baseline --no-e2e 2>gpurun_out/sweep_last.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('workers=$W batch=$B Q=$Q pairs/s=%.0f gcups=%.1f ms/step=%.1f shares=%s' % (d['value'], d['gcups'], d['ms_per_step'], {k: round(v,2) for k,v in r['share_of_step'].items()}))" || tail -3 gpurun_out/sweep_last.err
done
