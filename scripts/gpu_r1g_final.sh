#!/bin/bash
# Round-1 (session g) evidence on the committed build: bench lines, reference arm, launch lists, traffic, full ncu captures of the
# S0 camera-ray kernel (default workload) and of the pool closest-hit kernel (stand-in).
set -u
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
timeout 600 python bench.py --workload teapot-bench > gpurun_out/bench_teapot_bench.json 2> gpurun_out/bench_teapot_bench.err
timeout 600 python bench.py --workload cube-1080p --no-cpu-baseline > gpurun_out/bench_cube.json 2> gpurun_out/bench_cube.err
timeout 900 python bench.py --workload atrium-1080p --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_atrium.json 2> gpurun_out/bench_atrium.err
run() { # workload mode spp
  A="--workload $1 --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --mode $2"
  [ -n "${3:-}" ] && A="$A --spp $3"
  timeout 300 python bench.py $A > gpurun_out/plain_$1_$2.log 2>&1 && \
  timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/traffic_$1_$2.csv python bench.py $A > gpurun_out/ncu_traffic_$1_$2.log 2>&1
}
run teapot-1080p wavefront-hybrid-pool-s0
run atrium-1080p wavefront-hybrid-pool-s0 16
A="--steps 1 --warmup 0 --no-cpu-baseline --no-e2e --mode wavefront-hybrid-pool-s0"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"wf_primary" -s 1 -c 1 -o gpurun_out/prof_primary_s0 python bench.py $A > gpurun_out/ncu_primary.log 2>&1; echo "rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"wf_trace_pool" -s 6 -c 1 -o gpurun_out/prof_pool_atrium python bench.py $A --workload atrium-1080p --spp 16 > gpurun_out/ncu_pool.log 2>&1; echo "rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/bench_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["value"]), round(d["ms_per_step"],3), d.get("e2e",{}).get("value"), (d.get("roofline") or {}).get("frac"), d["config"].get("pipeline"))
    except Exception as e: print(f,"ERR",e)
PY
ls -la gpurun_out/*.ncu-rep
