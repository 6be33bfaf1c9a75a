#!/bin/bash
# Developer tool (GPU box): closing measurements of a session on ONE GPU: the GPU test suite, the driver's default bench
# line (C3) with its reference arm, and the ncu launch list of the same command.  TAG names the outputs.
TAG=${TAG:-r2o}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -n 2 gpurun_out/${TAG}_tests.log
timeout 900 python bench.py --steps ${STEPS:-5} --warmup 3 > gpurun_out/${TAG}_bench_c3.json 2> gpurun_out/${TAG}_bench_c3.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_c3_reference.json 2> gpurun_out/${TAG}_bench_c3_reference.err; echo "reference rc=$?"
B="--steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-isolated --no-traffic"
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 400 --csv --log-file gpurun_out/${TAG}_launches_c3.csv python bench.py $B > gpurun_out/${TAG}_ncu_c3.log 2>&1
python - <<PY
import json
for n in ("bench_c3", "bench_c3_reference"):
    try:
        d = json.loads(open("gpurun_out/${TAG}_%s.json" % n).read().strip().split("\n")[-1])
        e = d.get("e2e") or {}
        print(n, "pairs/s=%.0f e2e=%.0f ms/step=%.1f" % (d["value"], e.get("value", 0), d.get("ms_per_step", 0)))
    except Exception as ex:
        print(n, "no result", ex)
PY
