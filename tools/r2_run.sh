#!/bin/bash
# Developer tool (GPU box): tests + a few short bench lines of the state under test.  TAG names the outputs.
TAG=${TAG:-r2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/${TAG}_pytest.txt
cat gpurun_out/${TAG}_pytest.txt | tail -5
B="--no-cpu-baseline --no-e2e --no-isolated --no-traffic"
run() { # name, env, args
  env $2 timeout 600 python bench.py $3 > gpurun_out/${TAG}_$1.json 2> gpurun_out/${TAG}_$1.err || { echo "$1 FAILED"; tail -3 gpurun_out/${TAG}_$1.err; }
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_$1.json")); r=d["roofline"]
    print("[$1 | $2 | $3] pairs/s=%.0f ms/step=%.1f dp_ms/launch=%.3f live=%.3f refilled=%.0f shares=%s" % (d["value"], d["ms_per_step"], r["dp_ms_per_launch"] or 0, r["live_lane_fraction"] or 0, d["refilled_pairs_per_step"], {k: round(v,2) for k,v in r["share_of_step"].items()}))
except Exception as e:
    print("[$1] no result", e)
PY
}
run c3 "" "--steps 3 --warmup 1 $B"
run ew8 "" "--steps 2 --warmup 1 --emulate-world 8 $B"
run ew8_w1 "LARA_B200_WORKERS=1" "--steps 2 --warmup 1 --emulate-world 8 $B"
