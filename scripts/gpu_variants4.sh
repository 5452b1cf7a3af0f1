#!/bin/bash
set -u
mkdir -p gpurun_out
for v in default "$@"; do
  if [ $v = default ]; then unset MINIPATH_B200_LIB; else export MINIPATH_B200_LIB=$PWD/build/libminipath_b200_$v.so; fi
  timeout 600 python bench.py --workload atrium-1080p --spp 16 --no-cpu-baseline --no-e2e --steps 3 --mode wavefront-hybrid-pool > gpurun_out/v4_${v}_atrium.json 2> gpurun_out/v4_${v}_atrium.err
  timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 3 --mode wavefront-hybrid-pool > gpurun_out/v4_${v}_teapot.json 2> gpurun_out/v4_${v}_teapot.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/v4_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["value"]), round(d["ms_per_step"],2))
    except Exception as e: print(f,"ERR",open(f.replace(".json",".err")).read()[-300:])
PY
