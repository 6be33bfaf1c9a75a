# Developer tool (GPU box): GPU tests, then C3 / C5 sub-samples under the memory-limited refill.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
B="--no-cpu-baseline --no-e2e --no-isolated --no-traffic"
timeout 600 python bench.py --steps 2 --warmup 1 $B > gpurun_out/r2m_c3.json 2> gpurun_out/r2m_c3.err; grep -h "step:" gpurun_out/r2m_c3.err | tail -1
for n in 400 800; do
  LARA_B200_VERBOSE=${V:-0} timeout 600 python bench.py --workload c5 --nseq $n --steps 2 --warmup 1 $B > gpurun_out/r2m_c5_$n.json 2> gpurun_out/r2m_c5_$n.err
  grep -h "step:" gpurun_out/r2m_c5_$n.err | tail -1
done
LARA_B200_VERBOSE=1 timeout 600 python bench.py --workload c5 --nseq 800 --steps 1 --warmup 0 $B 2>&1 >/dev/null | grep "resident\|step:" | tail -40
