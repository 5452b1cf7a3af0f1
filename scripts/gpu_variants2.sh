#!/bin/bash
# A/B of library builds x node layouts on the stand-in (8 spp) + default workloads
set -u
mkdir -p gpurun_out
for v in default "$@"; do
  if [ $v = default ]; then unset MINIPATH_B200_LIB; else export MINIPATH_B200_LIB=$PWD/build/libminipath_b200_$v.so; fi
  for L in auto decoded; do
    timeout 600 python bench.py --workload atrium-1080p --spp 8 --no-cpu-baseline --no-e2e --steps 3 --layout $L > gpurun_out/var_${v}_$L.json 2> gpurun_out/var_${v}_$L.err
  done
done
unset MINIPATH_B200_LIB
timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 3 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
timeout 600 python bench.py --workload teapot-bench --no-cpu-baseline --no-e2e --steps 3 > gpurun_out/bench_teapot_bench.json 2> gpurun_out/bench_teapot_bench.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/var_*.json"))+["gpurun_out/bench_default.json","gpurun_out/bench_teapot_bench.json"]:
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["value"]), round(d["ms_per_step"],2), d["config"]["pipeline"], {k:round(v,1) for k,v in d["config"]["pipeline_ab_ms"].items()})
    except Exception as e: print(f,"ERR",open(f.replace(".json",".err")).read()[-300:])
PY
