#!/bin/bash
# HYBRID_POOL_S0 (depth-0 sun rays traced inside the camera-ray kernel): parity suite on all 11 contexts, smoke, bench lines
# (the A/B inside bench.py picks the pipeline), reference arm, traffic + launch list of the new pipeline.
set -u
mkdir -p gpurun_out
timeout 1000 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
timeout 600 python bench.py --workload atrium-1080p --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_atrium.json 2> gpurun_out/bench_atrium.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
for M in wavefront-hybrid-pool wavefront-hybrid-pool-s0; do
  A="--steps 3 --warmup 3 --no-cpu-baseline --no-e2e --mode $M"
  timeout 300 python bench.py $A > gpurun_out/ab_default_$M.json 2>/dev/null
  timeout 300 python bench.py $A --workload atrium-1080p --spp 32 > gpurun_out/ab_atrium_$M.json 2>/dev/null
done
run() { # workload mode spp
  A="--workload $1 --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --mode $2"
  [ -n "${3:-}" ] && A="$A --spp $3"
  timeout 300 python bench.py $A > gpurun_out/plain_$1_$2.log 2>&1 && \
  timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/traffic_$1_$2.csv python bench.py $A > gpurun_out/ncu_traffic_$1_$2.log 2>&1
}
run teapot-1080p wavefront-hybrid-pool-s0
run atrium-1080p wavefront-hybrid-pool-s0 16
tail -3 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/bench_*.json")+glob.glob("gpurun_out/ab_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["value"]), round(d["ms_per_step"],3), d.get("e2e",{}).get("value"), (d.get("roofline") or {}).get("frac"), d["config"].get("pipeline"), d["config"].get("pipeline_ab_ms"))
    except Exception as e: print(f,"ERR",e)
PY
