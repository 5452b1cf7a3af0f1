#!/bin/bash
# A/B of library builds on the S0 pipeline: default workload and the stand-in (32 spp)
set -u
mkdir -p gpurun_out
rm -f gpurun_out/v5_*
for v in default "$@"; do
  if [ $v = default ]; then unset MINIPATH_B200_LIB; else export MINIPATH_B200_LIB=$PWD/build/libminipath_b200_$v.so; fi
  A="--no-cpu-baseline --no-e2e --steps 4 --warmup 3 --mode wavefront-hybrid-pool-s0"
  timeout 300 python bench.py $A > gpurun_out/v5_${v}_default.json 2> gpurun_out/v5_${v}_default.err
  timeout 300 python bench.py $A --workload atrium-1080p --spp 32 > gpurun_out/v5_${v}_atrium.json 2> gpurun_out/v5_${v}_atrium.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/v5_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["value"]), round(d["ms_per_step"],3))
    except Exception as e: print(f,"ERR",open(f.replace(".json",".err")).read()[-300:])
PY
