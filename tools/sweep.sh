#!/bin/bash
# Developer tool (GPU box): bench.py over LARA_B200_WORKERS x LARA_B200_BATCH; prints one short line per run.
NSEQ=${NSEQ:-200}
for cfg in "$@"; do
  W=${cfg%%:*}; B=${cfg##*:}
  LARA_B200_WORKERS=$W LARA_B200_BATCH=$B timeout 600 python bench.py --nseq $NSEQ --steps 2 --warmup 1 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('workers=$W batch=$B pairs/s=%.0f gcups=%.1f ms/step=%.1f dp_kernel_gcups=%.0f shares=%s' % (d['value'], d['gcups'], d['ms_per_step'], r['dp_gcups_kernel_only'], {k: round(v,2) for k,v in r['share_of_step'].items()}))"
done
