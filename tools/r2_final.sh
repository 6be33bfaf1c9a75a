#!/bin/bash
# Developer tool (GPU box): the measurements of record of a round on ONE GPU: bench lines of C3 (the driver's default),
# C4, C2, C1 with their reference arms, ncu launch lists and one ncu --set full capture of C3 and C4.  TAG names the outputs.
TAG=${TAG:-r2}
mkdir -p gpurun_out
run() { # name, args
  timeout 1500 python bench.py $2 > gpurun_out/${TAG}_$1.json 2> gpurun_out/${TAG}_$1.err || { echo "$1 FAILED"; tail -3 gpurun_out/${TAG}_$1.err; }
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_$1.json").read().strip().split("\n")[-1])
    if d.get("impl") == "reference":
        print("[$1] reference %.1f pairs/s (%s)" % (d["value"], d["cpu_baseline"]["sample"][:60]))
    else:
        r=d["roofline"]; e=d.get("e2e") or {}; c=d.get("cpu_baseline") or {}
        print("[$1] pairs/s=%.0f e2e=%.0f ms/step=%.1f gcups_eff=%.0f computed=%.0f dp_ms/launch=%.3f frac=%.3f traffic=%s cpu=%s simd=%s same=%s" % (
            d["value"], e.get("value", 0), d["ms_per_step"], d["gcups_effective"], d["gcups_computed"], r["dp_ms_per_launch"] or 0, r["frac"] or 0,
            r.get("traffic"), c.get("value"), (c.get("simd_build") or {}).get("value"), c.get("output_identical_to_gpu_on_sample")))
except Exception as ex:
    print("[$1] no result", ex)
PY
}
run bench_c3 "--steps ${STEPS:-5} --warmup 3"
run bench_c3_reference "--impl reference --steps 2 --warmup 1"
run bench_c4 "--workload c4 --steps 5 --warmup 2"
run bench_c4_reference "--workload c4 --impl reference --steps 2 --warmup 1"
run bench_c2 "--workload c2 --steps 5 --warmup 3"
run bench_c1 "--workload c1 --steps 5 --warmup 3"
B="--steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-isolated --no-traffic"
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 400 --csv --log-file gpurun_out/${TAG}_launches_c3.csv python bench.py $B > gpurun_out/${TAG}_ncu_c3.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 400 --csv --log-file gpurun_out/${TAG}_launches_c4.csv python bench.py --workload c4 $B > gpurun_out/${TAG}_ncu_c4.log 2>&1
LARA_B200_WORKERS=1 ncu --set full --clock-control none --import-source on -k regex:"gotoh_|iterate_post" -s 60 -c 4 -o gpurun_out/${TAG}_full_c3 -f python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-isolated --no-traffic > gpurun_out/${TAG}_ncu2_c3.log 2>&1
LARA_B200_WORKERS=1 ncu --set full --clock-control none --import-source on -k regex:"gotoh_|iterate_post" -s 60 -c 4 -o gpurun_out/${TAG}_full_c4 -f python bench.py --workload c4 --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-isolated --no-traffic > gpurun_out/${TAG}_ncu2_c4.log 2>&1
ls -la gpurun_out | grep ${TAG}_ | head -40
