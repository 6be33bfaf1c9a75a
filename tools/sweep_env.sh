#!/bin/bash
# Developer tool (GPU box): bench.py once per argument, each argument a list of environment assignments
# ("LARA_B200_TUNE=1 LARA_B200_WORKERS=1"); prints one short line per run.
# usage: NSEQ=350 EXTRA="--emulate-world 8" tools/sweep_env.sh "" "A=1" ...
NSEQ=${NSEQ:-350}
for cfg in "$@"; do
  env $cfg timeout 900 python bench.py --nseq $NSEQ --steps ${STEPS:-1} --warmup 1 --no-cpu-baseline --no-e2e --no-isolated --no-traffic $EXTRA 2>gpurun_out/sweep_last.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('[%s %s] pairs/s=%.0f ms/step=%.1f dp_ms/launch=%.3f shares=%s' % ('$cfg', '$EXTRA', d['value'], d['ms_per_step'], r['dp_ms_per_launch'], {k: round(v,2) for k,v in r['share_of_step'].items()}))" || tail -3 gpurun_out/sweep_last.err
done
