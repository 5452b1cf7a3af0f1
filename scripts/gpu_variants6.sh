#!/bin/bash
# A/B of library builds (no default leg): S0 pipeline, default workload and the stand-in at 32 spp
set -u
mkdir -p gpurun_out
for v in "$@"; do
  export MINIPATH_B200_LIB=$PWD/build/libminipath_b200_$v.so
  A="--no-cpu-baseline --no-e2e --steps 3 --warmup 2 --mode wavefront-hybrid-pool-s0"
  timeout 200 python bench.py $A > gpurun_out/v6_${v}_default.json 2> gpurun_out/v6_${v}_default.err
  timeout 200 python bench.py $A --workload atrium-1080p --spp 32 > gpurun_out/v6_${v}_atrium.json 2> gpurun_out/v6_${v}_atrium.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/v6_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["value"]), round(d["ms_per_step"],3))
    except Exception as e: print(f,"ERR",open(f.replace(".json",".err")).read()[-300:])
PY
